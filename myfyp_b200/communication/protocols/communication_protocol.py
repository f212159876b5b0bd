"""The 13-method protocol contract (reference: communication/protocols/communication_protocol.py:27-198)."""

from __future__ import annotations

import abc
import functools
from typing import Any, Callable, Dict, List, Optional, Union

from myfyp_b200.communication.commands.command import Command
from myfyp_b200.communication.protocols.exceptions import ProtocolNotStartedError


def running(func):
    """Guard: the protocol's server must be running."""

    @functools.wraps(func)
    def wrapper(self, *args, **kwargs):
        if not self._server.is_running():
            raise ProtocolNotStartedError("The protocol has not been started.")
        return func(self, *args, **kwargs)

    return wrapper


class CommunicationProtocol(abc.ABC):
    """Transport-agnostic node-to-node communication."""

    @abc.abstractmethod
    def __init__(self, addr: Optional[str] = None, commands: Optional[List[Command]] = None) -> None: ...

    @abc.abstractmethod
    def start(self) -> None: ...

    @abc.abstractmethod
    def stop(self) -> None: ...

    @abc.abstractmethod
    def add_command(self, cmds: Union[Command, List[Command]]) -> None: ...

    @abc.abstractmethod
    def build_msg(self, cmd: str, args: Optional[List[str]] = None, round: Optional[int] = None) -> Any: ...

    @abc.abstractmethod
    def build_weights(self, cmd: str, round: int, serialized_model: Any, contributors: Optional[List[str]] = None, weight: int = 1) -> Any: ...

    @abc.abstractmethod
    def send(self, nei: str, msg: Any, raise_error: bool = False, remove_on_error: bool = True) -> None: ...

    @abc.abstractmethod
    def broadcast(self, msg: Any, node_list: Optional[List[str]] = None) -> None: ...

    @abc.abstractmethod
    def connect(self, addr: str, non_direct: bool = False) -> bool: ...

    @abc.abstractmethod
    def disconnect(self, nei: str, disconnect_msg: bool = True) -> None: ...

    @abc.abstractmethod
    def get_neighbors(self, only_direct: bool = False) -> Dict[str, Any]: ...

    @abc.abstractmethod
    def get_address(self) -> str: ...

    @abc.abstractmethod
    def wait_for_termination(self) -> None: ...

    @abc.abstractmethod
    def gossip_weights(
        self,
        early_stopping_fn: Callable[[], bool],
        get_candidates_fn: Callable[[], List[str]],
        status_fn: Callable[[], Any],
        model_fn: Callable[[str], Any],
        period: Optional[float] = None,
        create_connection: bool = False,
        wait_fn: Optional[Callable[[float], None]] = None,
    ) -> None: ...

    def broadcast_async(self, msg: Any, coalesce: bool = False, linger: float = 0.0) -> None:
        """Like :meth:`broadcast` but never blocks the caller on the peers (default: plain broadcast)."""
        self.broadcast(msg)

    # --- B200 data-plane hook (no counterpart in the reference) -------------------
    def fabric_reachable(self, nei: str) -> bool:
        """True when a device descriptor sent to ``nei`` can be dereferenced there."""
        return False
