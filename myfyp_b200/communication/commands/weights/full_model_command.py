"""``add_model`` = full aggregated model diffusion (reference: commands/weights/full_model_command.py:31-89)."""

import threading
from typing import Callable

from myfyp_b200.communication.commands.command import Command
from myfyp_b200.communication.commands.weights._common import bad_model_guard
from myfyp_b200.management.logger import logger


class FullModelCommand(Command):
    def __init__(self, state, stop: Callable[[], None], aggregator, learner, protocol=None) -> None:
        self.state, self.stop, self.aggregator, self.learner, self.protocol = state, stop, aggregator, learner, protocol
        self._mutex = threading.Lock()  # several neighbours may push the aggregate at the same instant

    @staticmethod
    def get_name() -> str:
        return "add_model"

    def adopt_early(self) -> bool:
        """Called by WaitAggregatedModelsStage once it waits: adopt an aggregate of this round that arrived before that."""
        st = self.state
        kept, st.early_full_model = st.early_full_model, None
        if kept is None or st.round is None or kept[0] != st.round:
            return False
        self.execute(kept[1], kept[0], weights=kept[2])
        return st.aggregated_model_event.is_set()

    def execute(self, source: str, round: int, weights=None, **kwargs) -> None:
        if weights is None:
            raise ValueError("Weights are required")
        st = self.state
        if st.round is None:
            logger.debug(st.addr, "❌ Tried to add a model while learning is not running")
            return
        if round == st.round + 1 and not st.rejoining:
            # the sender has already finished the round we are still voting in: keep its aggregate for when we wait for it
            st.early_full_model = (round, source, weights)
            return
        if round <= st.aggregate_round:
            return  # we hold this round's aggregate already: one of the several neighbours that push it at the same instant was first
        if round != st.round:
            if st.rejoining and round > st.round and st.experiment is not None:
                # we joined a running experiment believing it was in round st.round, but the network has moved on:
                # fast-forward to the round this aggregate belongs to
                logger.info(st.addr, f"⏩ Rejoining: fast-forward from round {st.round} to {round}.")
                st.experiment.round = round
            else:
                logger.debug(st.addr, f"Model reception in a late round ({round} != {st.round}).")
                return
        with bad_model_guard(st, self.stop, source, self.protocol):
            with self._mutex:
                if st.aggregated_model_event.is_set():
                    # Not waiting for it (yet).  A node outside the train set that is still tallying this round's votes - e.g.
                    # held up by the ballot of a peer that just died - will wait for exactly this model in a moment, and the
                    # pusher gives up after a few unanswered offers (reference: full_model_command.py:71-74 drops it).
                    st.early_full_model = (round, source, weights)
                    logger.debug(st.addr, "😲 Aggregated model not expected yet: kept.")
                    return
                logger.info(st.addr, "📦 Aggregated model received.")
                self.learner.set_model(weights)
                try:  # server-side state that travels with the model (Scaffold's c); never a reason to reject the model
                    self.aggregator.observe_global(self.learner.get_model())
                except Exception as e:  # noqa: BLE001
                    logger.debug(st.addr, f"observe_global failed: {e}")
                st.rejoining = False
                st.aggregate_round = round
                st.aggregated_model_event.set()
            st.notify()
