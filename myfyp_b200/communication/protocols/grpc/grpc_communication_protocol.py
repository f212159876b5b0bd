"""gRPC transport: unary RPCs (hand-written protobuf codec, optional mTLS) between hosts; nodes of the same process
short-circuit to a direct servicer call.  (Reference: communication/protocols/grpc/grpc_communication_protocol.py:49-263;
the facade itself is ``ComposedProtocol``.)"""

from __future__ import annotations

from typing import List, Optional

from myfyp_b200.communication.commands.command import Command
from myfyp_b200.communication.protocols.composed_protocol import ComposedProtocol
from myfyp_b200.communication.protocols.grpc.address import AddressParser
from myfyp_b200.communication.protocols.grpc.grpc_client import GrpcClient
from myfyp_b200.communication.protocols.grpc.grpc_neighbors import GrpcNeighbors
from myfyp_b200.communication.protocols.grpc.grpc_server import GrpcServer, local_server
from myfyp_b200.settings import Settings


class GrpcCommunicationProtocol(ComposedProtocol):
    def __init__(self, addr: str = "127.0.0.1", commands: Optional[List[Command]] = None) -> None:
        parsed = AddressParser(addr).get_parsed_address()
        self._assemble(parsed, GrpcNeighbors(parsed), GrpcClient, GrpcServer, commands)

    def fabric_reachable(self, nei: str) -> bool:
        """Same process (direct servicer call) or a rank of this box whose arena we have mapped."""
        if Settings.GRPC_LOCAL_SHORTCUT and local_server(nei) is not None:
            return True
        from myfyp_b200.parallel.ipc import address_on_fabric

        return address_on_fabric(nei)
