"""``partial_model`` (reference: commands/weights/partial_model_command.py:33-112).

``weights`` is either the reference's pickled byte blob or a device descriptor
(:class:`myfyp_b200.parallel.fabric.SlabRef`).  With a descriptor nothing is copied
here: the aggregator keeps the descriptor and the aggregation kernel pulls the
sender's slab out of (peer) HBM when it reduces.
"""

from typing import Callable, List, Optional

from myfyp_b200.communication.commands.command import Command
from myfyp_b200.communication.commands.message.models_aggregated_command import ModelsAggregatedCommand
from myfyp_b200.communication.commands.weights._common import bad_model_guard
from myfyp_b200.management.logger import logger
from myfyp_b200.management.trace import trace
from myfyp_b200.settings import Settings


class PartialModelCommand(Command):
    def __init__(self, state, stop: Callable[[], None], aggregator, comm_proto, learner) -> None:
        self.state, self.stop, self.aggregator = state, stop, aggregator
        self.communication_protocol, self.learner = comm_proto, learner

    @staticmethod
    def get_name() -> str:
        return "partial_model"

    def absorb_early(self) -> None:
        """Called by TrainStage once the round's train set is installed: file the offers that arrived one round early."""
        st = self.state
        with st.train_set_votes_lock:
            mine = [e for e in st.early_models if e[0] == st.round]
            st.early_models = [e for e in st.early_models if st.round is not None and e[0] > st.round]
        for _, source, weights, contributors, num_samples in mine:
            self.absorb(source, weights, contributors, num_samples)

    def execute(self, source: str, round: int, weights=None, contributors: Optional[List[str]] = None, num_samples: Optional[int] = None, **kwargs) -> None:
        if weights is None or contributors is None or num_samples is None:
            raise ValueError("Weights, contributors and weight are required")
        st = self.state
        if st.round is None:
            logger.debug(st.addr, "Tried to add a model while learning is not running")
            return
        if round == st.round + 1 or (round == st.round and not self.aggregator.is_collecting() and st.addr not in self.aggregator.get_aggregated_models()):
            # The sender is ahead of us: a round ahead, or already training while we still tally this round's votes (we are still finishing the previous round, e.g. waiting for the ballot of a peer
            # that died).  Its published snapshot is immutable for the round, so the offer is kept and filed when we get
            # there; dropping it (reference: partial_model_command.py:66-70) makes the sender retry for a few ticks, give up
            # and leave the round - after which this node waits for a model that never comes until AGGREGATION_TIMEOUT.
            with st.train_set_votes_lock:
                st.early_models = [e for e in st.early_models if e[0] >= st.round][-64:] + [(round, source, weights, list(contributors), num_samples)]
            return
        if round != st.round:
            logger.debug(st.addr, f"Model reception in a late round ({round} != {st.round}).")
            return
        if len(st.train_set) == 0:
            logger.error(st.addr, "Model Reception when there is no trainset")
            return
        self.absorb(source, weights, contributors, num_samples)

    def absorb(self, source: str, weights, contributors: List[str], num_samples: int) -> None:
        """File one offered model with the aggregator and announce what we hold now."""
        st = self.state
        trace(st.addr, f"recv_model<{source} {len(contributors)}")
        with bad_model_guard(st, self.stop, source, self.communication_protocol):
            model = self.learner.get_model().build_copy(params=weights, num_samples=num_samples, contributors=list(contributors))
            models_added = self.aggregator.add_model(model)
            trace(st.addr, f"added<{source} held={len(models_added)}")
            # the sender evidently holds what it sent: no need to wait for its announcement to know that
            held = st.models_aggregated.get(source, [])
            st.models_aggregated[source] = held + [c for c in contributors if c not in held]
            if models_added != []:
                # we are inside the sender's RPC / call stack: announce from the gossiper thread, never block here
                # (a synchronous fan-out from inside a handler can exhaust the peers' RPC workers)
                # announcements are cumulative (receivers merge them), so a newer one supersedes an unsent older one
                # and a burst of arrivals is worth one announcement: only the complete set goes out without delay
                complete = len(set(models_added)) >= len(st.train_set)
                self.communication_protocol.broadcast_async(
                    self.communication_protocol.build_msg(ModelsAggregatedCommand.get_name(), models_added, round=st.round), coalesce=True,
                    linger=0.0 if complete else Settings.ANNOUNCE_LINGER,
                )
