"""``models_ready``: "I hold the aggregated model of round r" (reference: commands/message/models_ready_command.py:26-62).

The reference stores the RECEIVER's round for a message of its current or previous round.  That is wrong by one
exactly when it matters: a peer's announcement for round r that arrives after this node has moved on to r + 1 is filed as
"has the model of r + 1", the peer is never offered that model in ``GossipModelStage``, and - when this node is its only route
to the train set - it sits in ``WaitAggregatedModelsStage`` until ``AGGREGATION_TIMEOUT`` while everybody else waits for its
ballot.  Here the announced round itself is stored (never moving backwards), so a peer that is behind stays visibly behind.
"""

from myfyp_b200.communication.commands.command import Command
from myfyp_b200.management.logger import logger


class ModelsReadyCommand(Command):
    def __init__(self, state) -> None:
        self.state = state

    @staticmethod
    def get_name() -> str:
        return "models_ready"

    def execute(self, source: str, round: int, *args, **kwargs) -> None:
        st = self.state
        if st.round is None:
            return
        if round < st.round - 1:
            logger.debug(st.addr, f"models_ready from {source} for round {round}, long past ({st.round}). Ignored.")
            return
        if round > st.nei_status.get(source, -2):
            st.nei_status[source] = round
            # only the status of DIRECT neighbours is waited for (GossipModelStage pushes the aggregate to those that are behind)
            probe = st.is_direct
            if probe is None or probe(source):
                st.notify()
