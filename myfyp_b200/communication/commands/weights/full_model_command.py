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

    def execute(self, source: str, round: int, weights=None, **kwargs) -> None:
        if weights is None:
            raise ValueError("Weights are required")
        st = self.state
        if st.round is None:
            logger.debug(st.addr, "❌ Tried to add a model while learning is not running")
            return
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
                    logger.debug(st.addr, "😲 Aggregated model not expected.")
                    return
                logger.info(st.addr, "📦 Aggregated model received.")
                self.learner.set_model(weights)
                try:  # server-side state that travels with the model (Scaffold's c); never a reason to reject the model
                    self.aggregator.observe_global(self.learner.get_model())
                except Exception as e:  # noqa: BLE001
                    logger.debug(st.addr, f"observe_global failed: {e}")
                st.rejoining = False
                st.aggregated_model_event.set()
            st.notify()
