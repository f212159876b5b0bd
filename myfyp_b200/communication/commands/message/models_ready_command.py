"""``models_ready`` (reference: commands/message/models_ready_command.py:26-62).

Like the reference it stores the *receiver's* round.
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
        if round in (st.round - 1, st.round):
            st.nei_status[source] = st.round
            st.notify()
        else:
            logger.debug(st.addr, f"models_ready from {source} in a late round ({round} != {st.round}). Ignored.")
