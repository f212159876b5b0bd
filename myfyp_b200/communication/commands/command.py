"""Command ABC (reference: p2pfl/communication/commands/command.py:24-43)."""

from __future__ import annotations

import abc


class Command(abc.ABC):
    """An RPC handler: a wire-visible name plus an ``execute``."""

    @staticmethod
    @abc.abstractmethod
    def get_name() -> str:
        """Wire-visible command name."""

    @abc.abstractmethod
    def execute(self, source: str, round: int, *args, **kwargs) -> None:
        """Run the command for a message coming from ``source`` at ``round``."""
