"""Stage ABC and early-stop helper (reference: p2pfl/stages/stage.py:26-66)."""

from __future__ import annotations

import abc
from typing import Type, Union


class EarlyStopException(Exception):
    """Learning was interrupted (``state.round`` became None)."""


def check_early_stop(state, raise_exception: bool = True) -> bool:
    if state.round is None:
        if raise_exception:
            raise EarlyStopException("Process stopped early")
        return True
    return False


class Stage(abc.ABC):
    @staticmethod
    @abc.abstractmethod
    def name() -> str: ...

    @staticmethod
    @abc.abstractmethod
    def execute(**kwargs) -> Union[Type["Stage"], None]: ...
