"""gRPC protocol = address parser + neighbours + client + gossiper + server + heartbeater
(reference: communication/protocols/grpc/grpc_communication_protocol.py:49-263)."""

from __future__ import annotations

from typing import Any, Callable, Dict, List, Optional, Union

from myfyp_b200.communication.commands.command import Command
from myfyp_b200.communication.commands.message.heartbeat_command import HeartbeatCommand
from myfyp_b200.communication.protocols.communication_protocol import CommunicationProtocol, running
from myfyp_b200.communication.protocols.gossiper import Gossiper
from myfyp_b200.communication.protocols.grpc.address import AddressParser
from myfyp_b200.communication.protocols.grpc.grpc_client import GrpcClient
from myfyp_b200.communication.protocols.grpc.grpc_neighbors import GrpcNeighbors
from myfyp_b200.communication.protocols.grpc.grpc_server import GrpcServer, local_server
from myfyp_b200.communication.protocols.heartbeater import Heartbeater
from myfyp_b200.settings import Settings


class GrpcCommunicationProtocol(CommunicationProtocol):
    def __init__(self, addr: str = "127.0.0.1", commands: Optional[List[Command]] = None) -> None:
        self.addr = AddressParser(addr).get_parsed_address()
        self._neighbors = GrpcNeighbors(self.addr)
        self._client = GrpcClient(self.addr, self._neighbors)
        self._gossiper = Gossiper(self.addr, self._client)
        self._server = GrpcServer(self.addr, self._gossiper, self._neighbors, commands)
        self._heartbeater = Heartbeater(self.addr, self._neighbors, self._client)
        self._server.add_command(HeartbeatCommand(self._heartbeater))
        if commands:
            self._server.add_command(commands)

    def get_address(self) -> str:
        return self.addr

    def start(self) -> None:
        self._server.start()
        self._heartbeater.start()
        self._gossiper.start()

    @running
    def stop(self) -> None:
        self._heartbeater.stop()
        self._gossiper.stop()
        self._neighbors.clear_neighbors()
        self._server.stop()

    def add_command(self, cmds: Union[Command, List[Command]]) -> None:
        self._server.add_command(cmds)

    @running
    def connect(self, addr: str, non_direct: bool = False) -> bool:
        return self._neighbors.add(addr, non_direct=non_direct)

    @running
    def disconnect(self, nei: str, disconnect_msg: bool = True) -> None:
        self._neighbors.remove(nei, disconnect_msg=disconnect_msg)

    def build_msg(self, cmd: str, args: Optional[List[str]] = None, round: Optional[int] = None) -> Any:
        return self._client.build_message(cmd, args, round)

    def build_weights(self, cmd: str, round: int, serialized_model: Any, contributors: Optional[List[str]] = None, weight: int = 1) -> Any:
        return self._client.build_weights(cmd, round, serialized_model, contributors, weight)

    @running
    def send(self, nei: str, msg: Any, raise_error: bool = False, remove_on_error: bool = True) -> None:
        self._client.send(nei, msg, raise_error=raise_error, remove_on_error=remove_on_error)

    @running
    def broadcast(self, msg: Any, node_list: Optional[List[str]] = None) -> None:
        if isinstance(msg, dict) and "hash" in msg:
            self._gossiper.check_and_set_processed(msg["hash"])
        self._client.broadcast(msg, node_list)

    @running
    def broadcast_async(self, msg: Any, coalesce: bool = False, linger: float = 0.0) -> None:
        """Queue ``msg`` for every direct neighbour on the gossiper thread and return at once (used inside handlers).
        ``coalesce``: a queued, not yet sent message of the same command and round is superseded by this one;
        ``linger``: hold a coalescable message back this long so that a burst of them collapses into one."""
        if isinstance(msg, dict) and "hash" in msg:
            self._gossiper.check_and_set_processed(msg["hash"])
        nodes = list(self._neighbors.get_all(only_direct=True))
        if isinstance(msg, dict) and "hash" in msg:
            msg["seen"] = [self.addr, *nodes]
        key = (msg.get("cmd"), msg.get("round")) if coalesce and isinstance(msg, dict) else None
        self._gossiper.add_message(msg, nodes, coalesce_key=key, linger=linger)

    @running
    def get_neighbors(self, only_direct: bool = False) -> Dict[str, Any]:
        return self._neighbors.get_all(only_direct)

    @running
    def wait_for_termination(self) -> None:
        self._server.wait_for_termination()

    @running
    def gossip_weights(
        self,
        early_stopping_fn: Callable[[], bool],
        get_candidates_fn: Callable[[], List[str]],
        status_fn: Callable[[], Any],
        model_fn: Callable[[str], Any],
        period: Optional[float] = None,
        create_connection: bool = False,
        wait_fn: Optional[Callable[[float], None]] = None,
    ) -> None:
        if period is None:
            period = Settings.GOSSIP_MODELS_PERIOD
        self._gossiper.gossip_weights(early_stopping_fn, get_candidates_fn, status_fn, model_fn, period, create_connection, wait_fn)

    def on_neighbors_changed(self, fn: Callable[[], None]) -> None:
        self._neighbors.on_change(fn)

    def fabric_reachable(self, nei: str) -> bool:
        """Same process (direct servicer call) or a rank of this box whose arena we have mapped."""
        if Settings.GRPC_LOCAL_SHORTCUT and local_server(nei) is not None:
            return True
        from myfyp_b200.parallel.ipc import address_on_fabric

        return address_on_fabric(nei)
