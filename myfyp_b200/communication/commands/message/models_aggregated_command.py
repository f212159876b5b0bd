"""``models_aggregated`` (reference: commands/message/models_agregated_command.py:26-56)."""

from myfyp_b200.communication.commands.command import Command
from myfyp_b200.management.logger import logger


class ModelsAggregatedCommand(Command):
    def __init__(self, state) -> None:
        self.state = state

    @staticmethod
    def get_name() -> str:
        return "models_aggregated"

    def execute(self, source: str, round: int, *args, **kwargs) -> None:
        if round == self.state.round:
            self.state.models_aggregated[source] = list(args)
            self.state.notify()
        else:
            logger.debug(self.state.addr, f"models_aggregated from {source} in a late round ({round} != {self.state.round}). Ignored.")
