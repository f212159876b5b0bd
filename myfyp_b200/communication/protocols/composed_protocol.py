"""One protocol facade for every transport.

A transport (in-memory, gRPC, unix socket) contributes five parts - a neighbour table, a client, a server, and the two
transport-agnostic threads (``Gossiper``, ``Heartbeater``) wired to them - by calling :meth:`_assemble` from its constructor;
everything the rest of the framework calls (the 13 methods of the reference's ``CommunicationProtocol`` contract, plus
``broadcast_async`` / ``on_neighbors_changed`` / ``fabric_reachable``) is implemented once, here, on top of those parts.
(The reference writes the facade out per transport: memory_communication_protocol.py:64-268, grpc_communication_protocol.py:49-263.)
"""

from __future__ import annotations

from typing import Any, Callable, Dict, List, Optional, Union

from myfyp_b200.communication.commands.command import Command
from myfyp_b200.communication.commands.message.heartbeat_command import HeartbeatCommand
from myfyp_b200.communication.commands.message.node_down_command import NodeDownCommand
from myfyp_b200.communication.protocols.communication_protocol import CommunicationProtocol, running
from myfyp_b200.communication.protocols.gossiper import Gossiper
from myfyp_b200.communication.protocols.heartbeater import Heartbeater
from myfyp_b200.settings import Settings


def _is_flood(msg: Any) -> bool:
    return isinstance(msg, dict) and "hash" in msg


class ComposedProtocol(CommunicationProtocol):
    addr: str

    def __init__(self, addr: Optional[str] = None, commands: Optional[List[Command]] = None) -> None:  # pragma: no cover - transports override
        raise NotImplementedError("instantiate a transport (InMemory / Grpc / Uds CommunicationProtocol)")

    def _assemble(self, addr: str, neighbors, client_cls, server_cls, commands: Optional[List[Command]]) -> None:
        self.addr = addr
        self._neighbors = neighbors
        self._client = client_cls(addr, neighbors)
        self._gossiper = Gossiper(addr, self._client)
        self._server = server_cls(addr, self._gossiper, neighbors, commands)
        self._heartbeater = Heartbeater(addr, neighbors, self._client)
        self._server.add_command(HeartbeatCommand(self._heartbeater))
        self._server.add_command(NodeDownCommand(addr, neighbors))
        self._announced_lost: Dict[str, float] = {}
        neighbors.on_lost(self._announce_lost)
        if commands:
            self._server.add_command(commands)

    # ------------------------------------------------------------------ life cycle
    def get_address(self) -> str:
        return self.addr

    def start(self) -> None:
        self._server.start()
        self._heartbeater.start()
        self._gossiper.start()

    @running
    def stop(self) -> None:
        if Settings.FAILURE_DISSEMINATION:  # goodbye: nodes that know us only through the flood need not wait for our beats to time out
            try:
                self._client.broadcast(self.build_msg(NodeDownCommand.get_name(), [self.addr]))
            except Exception:  # noqa: BLE001
                pass
        self._heartbeater.stop()
        self._gossiper.stop()
        self._neighbors.clear_neighbors()
        self._server.stop()

    @running
    def wait_for_termination(self) -> None:
        self._server.wait_for_termination()

    def add_command(self, cmds: Union[Command, List[Command]]) -> None:
        self._server.add_command(cmds)

    # ------------------------------------------------------------------ neighbours
    @running
    def connect(self, addr: str, non_direct: bool = False) -> bool:
        return self._neighbors.add(addr, non_direct=non_direct)

    @running
    def disconnect(self, nei: str, disconnect_msg: bool = True) -> None:
        self._neighbors.remove(nei, disconnect_msg=disconnect_msg)

    @running
    def get_neighbors(self, only_direct: bool = False) -> Dict[str, Any]:
        return self._neighbors.get_all(only_direct)

    def on_neighbors_changed(self, fn: Callable[[], None]) -> None:
        self._neighbors.on_change(fn)

    def _announce_lost(self, dead: str) -> None:
        """A direct neighbour of ours is gone: tell everybody (``node_down``), once per few seconds and address."""
        import time as _t

        if not Settings.FAILURE_DISSEMINATION:
            return
        now = _t.monotonic()
        if now - self._announced_lost.get(dead, -1e9) < 2.0:
            return
        self._announced_lost[dead] = now
        try:
            self.broadcast_async(self.build_msg(NodeDownCommand.get_name(), [dead]))
        except Exception:  # noqa: BLE001 - protocol stopping
            pass

    # ------------------------------------------------------------------ messages
    def build_msg(self, cmd: str, args: Optional[List[str]] = None, round: Optional[int] = None) -> Any:
        return self._client.build_message(cmd, args, round)

    def build_weights(self, cmd: str, round: int, serialized_model: Any, contributors: Optional[List[str]] = None, weight: int = 1) -> Any:
        return self._client.build_weights(cmd, round, serialized_model, contributors, weight)

    @running
    def send(self, nei: str, msg: Any, raise_error: bool = False, remove_on_error: bool = True) -> None:
        self._client.send(nei, msg, raise_error=raise_error, remove_on_error=remove_on_error)

    @running
    def broadcast(self, msg: Any, node_list: Optional[List[str]] = None) -> None:
        if _is_flood(msg):  # the originator has "seen" its own message: the flood never comes back to it
            self._gossiper.check_and_set_processed(msg["hash"])
        self._client.broadcast(msg, node_list)

    @running
    def broadcast_async(self, msg: Any, coalesce: bool = False, linger: float = 0.0) -> None:
        """Queue ``msg`` for every direct neighbour on the gossiper thread and return at once (used inside handlers).
        ``coalesce``: a queued, not yet sent message of the same command and round is superseded by this one;
        ``linger``: hold a coalescable message back this long so that a burst of them collapses into one."""
        nodes = list(self._neighbors.direct_addrs())
        if _is_flood(msg):
            self._gossiper.check_and_set_processed(msg["hash"])
            msg["seen"] = [self.addr, *nodes]
        key = (self.addr, msg.get("cmd"), msg.get("round")) if coalesce and isinstance(msg, dict) else None  # own messages only
        self._gossiper.add_message(msg, nodes, coalesce_key=key, linger=linger)

    @running
    def gossip_weights(self, early_stopping_fn: Callable[[], bool], get_candidates_fn: Callable[[], List[str]], status_fn: Callable[[], Any],
                       model_fn: Callable[[str], Any], period: Optional[float] = None, create_connection: bool = False,
                       wait_fn: Optional[Callable[[float], None]] = None) -> None:
        self._gossiper.gossip_weights(early_stopping_fn, get_candidates_fn, status_fn, model_fn,
                                      Settings.GOSSIP_MODELS_PERIOD if period is None else period, create_connection, wait_fn)
