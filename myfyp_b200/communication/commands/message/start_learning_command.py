"""``start_learning`` (reference: commands/message/start_learning_command.py:26-60).

Extension (rejoin): an optional third argument carries the round the
sender is currently in, so a node that (re)joins a running experiment enters at that
round instead of round 0.
"""

from typing import Callable

from myfyp_b200.communication.commands.command import Command


class StartLearningCommand(Command):
    def __init__(self, start_learning_fn: Callable[..., None]) -> None:
        self.__start_learning_fn = start_learning_fn

    @staticmethod
    def get_name() -> str:
        return "start_learning"

    def execute(self, source: str, round: int, *args, **kwargs) -> None:
        if len(args) < 2:
            raise ValueError("Learning rounds and epochs are required")
        rounds, epochs = int(args[0]), int(args[1])
        if len(args) >= 3:
            self.__start_learning_fn(rounds, epochs, start_round=int(args[2]))
        else:
            self.__start_learning_fn(rounds, epochs)
