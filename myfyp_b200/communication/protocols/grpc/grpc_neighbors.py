"""gRPC neighbour table (reference: communication/protocols/grpc/grpc_neighbors.py:35-144).

direct neighbour = ``(channel, stub, last_beat)`` (+ handshake RPC); non-direct = ``(None, None, last_beat)``.
A *stub* here is a tiny object with ``handshake/disconnect/send`` callables - bound either to a channel or,
when the peer lives in this process, straight to its servicer.
"""

from __future__ import annotations

import os
import time as _time
from typing import Any, Optional, Tuple

import grpc

from myfyp_b200.communication.protocols.grpc import wire
from myfyp_b200.communication.protocols.grpc.grpc_server import MAX_MSG, SERVICE, local_server
from myfyp_b200.communication.protocols.neighbors import Neighbors
from myfyp_b200.management.logger import logger
from myfyp_b200.settings import Settings


class ChannelStub:
    def __init__(self, channel: grpc.Channel) -> None:
        self._handshake = channel.unary_unary(f"/{SERVICE}/handshake", request_serializer=wire.encode_handshake, response_deserializer=wire.decode_response)
        self._disconnect = channel.unary_unary(f"/{SERVICE}/disconnect", request_serializer=wire.encode_handshake, response_deserializer=wire.decode_empty)
        self._send = channel.unary_unary(f"/{SERVICE}/send", request_serializer=wire.encode_root, response_deserializer=wire.decode_response)
        self.local = False

    def handshake(self, addr: str, timeout: float) -> Optional[str]:
        return self._handshake(addr, timeout=timeout)

    def disconnect(self, addr: str, timeout: float) -> None:
        self._disconnect(addr, timeout=timeout)

    def send(self, msg, timeout: float) -> Optional[str]:
        return self._send(msg, timeout=timeout)


class LocalStub:
    """Peer in this process: call its servicer directly (descriptors stay Python objects)."""

    def __init__(self, addr: str) -> None:
        self._addr = addr
        self.local = True

    def _srv(self):
        srv = local_server(self._addr)
        if srv is None:
            raise grpc.RpcError(f"{self._addr} is gone")
        return srv

    def handshake(self, addr: str, timeout: float) -> Optional[str]:
        return self._srv().handshake(addr)

    def disconnect(self, addr: str, timeout: float) -> None:
        self._srv().disconnect(addr)

    def send(self, msg, timeout: float) -> Optional[str]:
        return self._srv().send(dict(msg))


def open_channel(addr: str) -> grpc.Channel:
    opts = [("grpc.max_send_message_length", MAX_MSG), ("grpc.max_receive_message_length", MAX_MSG)]
    if Settings.USE_SSL and os.path.isfile(Settings.CLIENT_KEY) and os.path.isfile(Settings.CLIENT_CRT):
        with open(Settings.CLIENT_KEY, "rb") as k, open(Settings.CLIENT_CRT, "rb") as c, open(Settings.CA_CRT, "rb") as ca:
            creds = grpc.ssl_channel_credentials(root_certificates=ca.read(), private_key=k.read(), certificate_chain=c.read())
        return grpc.secure_channel(addr, creds, options=opts)
    return grpc.insecure_channel(addr, options=opts)


def make_stub(addr: str) -> Tuple[Optional[grpc.Channel], Any]:
    if Settings.GRPC_LOCAL_SHORTCUT and local_server(addr) is not None:
        return None, LocalStub(addr)
    ch = open_channel(addr)
    return ch, ChannelStub(ch)


class GrpcNeighbors(Neighbors):
    def local_server(self, addr: str):
        """The peer's server object when it lives in this process (``LocalStub``), else None."""
        return local_server(addr)

    def make_stub(self, addr: str) -> Tuple[Any, Any]:
        """``(closable channel or None, stub)`` for ``addr`` - the transport hook subclasses override."""
        return make_stub(addr)

    def refresh_or_add(self, addr: str, time: float) -> None:
        with self.neis_lock:
            cur = self.neis.get(addr)
            if cur is not None:
                self.neis[addr] = (cur[0], cur[1], time)
                return
        self.add(addr, non_direct=True)

    def connect(self, addr: str, non_direct: bool = False, handshake_msg: bool = True) -> Tuple[Any, Any, float]:
        if non_direct:
            logger.debug(self.self_addr, f"🔍 Found node {addr}")
            return (None, None, _time.time())
        logger.debug(self.self_addr, f"🤝 Adding {addr}")
        channel, stub = self.make_stub(addr)
        try:
            if handshake_msg:
                err = stub.handshake(self.self_addr, Settings.GRPC_TIMEOUT)
                if err:
                    raise Exception(f"Cannot add a neighbor: {err}")
            return (channel, stub, _time.time())
        except Exception:
            if channel is not None:
                channel.close()
            raise

    def disconnect(self, addr: str, disconnect_msg: bool = True, entry: Any = None) -> None:
        if entry is None:
            entry = self.neis.get(addr)
        if entry is None:
            return
        channel, stub, _ = entry
        try:
            if disconnect_msg and stub is not None:
                stub.disconnect(self.self_addr, Settings.GRPC_TIMEOUT)
        except Exception:  # noqa: BLE001
            pass
        finally:
            if channel is not None:
                channel.close()
