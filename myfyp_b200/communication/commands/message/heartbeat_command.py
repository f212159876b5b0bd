"""``beat`` (reference: commands/message/heartbeat_command.py:27-52)."""

from myfyp_b200.communication.commands.command import Command


class HeartbeatCommand(Command):
    def __init__(self, heartbeat) -> None:
        self.__heartbeat = heartbeat

    @staticmethod
    def get_name() -> str:
        return "beat"

    def execute(self, source: str, round: int, *args, **kwargs) -> None:
        if len(args) < 1:
            raise ValueError("Time is required")
        self.__heartbeat.beat(source, time=float(args[0]))
