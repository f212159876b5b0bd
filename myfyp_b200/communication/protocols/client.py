"""Client ABC (reference: communication/protocols/client.py:25-89)."""

from __future__ import annotations

import abc
from typing import Any, List, Optional


class Client(abc.ABC):
    @abc.abstractmethod
    def build_message(self, cmd: str, args: Optional[List[str]] = None, round: Optional[int] = None) -> Any: ...

    @abc.abstractmethod
    def build_weights(self, cmd: str, round: int, serialized_model: Any, contributors: Optional[List[str]] = None, weight: int = 1) -> Any: ...

    @abc.abstractmethod
    def send(self, nei: str, msg: Any, create_connection: bool = False, raise_error: bool = False, remove_on_error: bool = True) -> None: ...

    @abc.abstractmethod
    def broadcast(self, msg: Any, node_list: Optional[List[str]] = None) -> None: ...
