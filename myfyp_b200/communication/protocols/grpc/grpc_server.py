"""gRPC server (reference: communication/protocols/grpc/grpc_server.py:44-237).

Three unary RPCs under ``/node.NodeServices/``: ``handshake``, ``disconnect``, ``send``; 1 GiB message cap;
optional mutual TLS from the ``Settings`` certificate paths.  Built on grpc's generic handlers with the
hand-written codec of :mod:`wire`.  The worker pool is larger than the reference's 2 so a big
model transfer cannot starve heartbeats.  Servers also register in a process-local table so that peers
living in the same process can call the servicer directly (no socket, no serialisation).
"""

from __future__ import annotations

import os
import threading
import time
from concurrent import futures
from typing import Any, Dict, List, Optional, Union

import grpc

from myfyp_b200.communication.commands.command import Command
from myfyp_b200.communication.protocols.gossiper import Gossiper
from myfyp_b200.communication.protocols.grpc import wire
from myfyp_b200.management.logger import logger
from myfyp_b200.settings import Settings

MAX_MSG = 1024 * 1024 * 1024
SERVICE = "node.NodeServices"

_local_servers: Dict[str, "GrpcServer"] = {}
_local_lock = threading.Lock()


def local_server(addr: str) -> Optional["GrpcServer"]:
    return _local_servers.get(addr)


class GrpcServer:
    def __init__(self, addr: str, gossiper: Gossiper, neighbors, commands: Optional[List[Command]] = None) -> None:
        self.addr = addr
        self._commands: Dict[str, Command] = {c.get_name(): c for c in (commands or [])}
        self._gossiper = gossiper
        self._neighbors = neighbors
        self._server: Optional[grpc.Server] = None
        self._started = False
        self._terminated = threading.Event()

    @property
    def gossiper(self) -> Gossiper:
        return self._gossiper

    # ------------------------------------------------------------------ lifecycle
    def start(self, wait: bool = False) -> None:
        opts = [("grpc.max_send_message_length", MAX_MSG), ("grpc.max_receive_message_length", MAX_MSG), ("grpc.so_reuseport", 0)]
        self._server = grpc.server(futures.ThreadPoolExecutor(max_workers=32, thread_name_prefix=f"grpc-{self.addr}"), options=opts)
        handlers = {
            "handshake": grpc.unary_unary_rpc_method_handler(self._rpc_handshake, request_deserializer=wire.decode_handshake, response_serializer=wire.encode_response),
            "disconnect": grpc.unary_unary_rpc_method_handler(self._rpc_disconnect, request_deserializer=wire.decode_handshake, response_serializer=wire.encode_empty),
            "send": grpc.unary_unary_rpc_method_handler(self._rpc_send, request_deserializer=wire.decode_root, response_serializer=wire.encode_response),
        }
        self._server.add_generic_rpc_handlers((grpc.method_handlers_generic_handler(SERVICE, handlers),))
        creds = self._server_credentials()
        port = self._server.add_secure_port(self.addr, creds) if creds is not None else self._server.add_insecure_port(self.addr)
        if port == 0 and not self.addr.startswith("unix:"):
            raise RuntimeError(f"cannot bind {self.addr}")
        self._server.start()
        self._started = True
        self._terminated.clear()
        with _local_lock:
            _local_servers[self.addr] = self
        logger.debug(self.addr, f"gRPC started at {self.addr}")
        if wait:
            self.wait_for_termination()

    @staticmethod
    def _server_credentials():
        if Settings.USE_SSL and os.path.isfile(Settings.SERVER_KEY) and os.path.isfile(Settings.SERVER_CRT):
            with open(Settings.SERVER_KEY, "rb") as k, open(Settings.SERVER_CRT, "rb") as c, open(Settings.CA_CRT, "rb") as ca:
                return grpc.ssl_server_credentials([(k.read(), c.read())], root_certificates=ca.read(), require_client_auth=True)
        return None

    def stop(self) -> None:
        with _local_lock:
            if _local_servers.get(self.addr) is self:
                del _local_servers[self.addr]
        self._started = False
        if self._server is not None:
            self._server.stop(0)
            self._server = None
        self._terminated.set()

    def wait_for_termination(self) -> None:
        self._terminated.wait()

    def is_running(self) -> bool:
        return self._started

    def add_command(self, cmds: Union[Command, List[Command]]) -> None:
        for c in cmds if isinstance(cmds, list) else [cmds]:
            self._commands[c.get_name()] = c

    # ------------------------------------------------------------------ servicer logic (also called in-process)
    def handshake(self, addr: str) -> Optional[str]:
        if not self._started:
            return "server not running"
        if self._neighbors.add(addr, non_direct=False, handshake_msg=False) or self._neighbors.exists(addr):
            return None
        return "Cannot add the node (duplicated or wrong direction)"

    def disconnect(self, addr: str) -> None:
        self._neighbors.remove(addr, disconnect_msg=False)

    def send(self, msg: Dict[str, Any]) -> Optional[str]:
        """Dedupe -> dispatch -> TTL re-gossip (order of the reference's gRPC server, grpc_server.py:161-217)."""
        if not self._started:
            return "server not running"
        is_weights = "weight" in msg
        if not is_weights and not self._gossiper.check_and_set_processed(msg["hash"]):
            return None
        cmd = msg["cmd"]
        if cmd != "beat":
            logger.log_bytes("msgs/" + cmd, 1)  # control-plane message census (benchmarks report it per round)
        if not is_weights and cmd != "beat" and msg["source"] != self.addr:
            # any message proves its origin is alive: learn about it now rather than at its next heartbeat
            self._neighbors.refresh_or_add(msg["source"], time.time())
        if logger.is_debug() and (cmd != "beat" or not Settings.EXCLUDE_BEAT_LOGS):
            logger.debug(self.addr, f"📫 {cmd.upper()} received from {msg['source']}")
        handler = self._commands.get(cmd)
        if handler is None:
            logger.error(self.addr, f"Unknown command: {cmd} from {msg['source']}")
            return f"Unknown command: {cmd}"
        try:
            if is_weights:
                handler.execute(msg["source"], msg["round"], weights=_maybe_descriptor(msg["weights"]), contributors=msg["contributors"], num_samples=msg["weight"])
            else:
                handler.execute(msg["source"], msg["round"], *msg["args"])
        except Exception as e:  # noqa: BLE001
            text = f"Error while processing command: {cmd}: {e}"
            logger.error(self.addr, text)
            return text
        if not is_weights and msg["ttl"] > 0:
            fwd = dict(msg)
            fwd["ttl"] = msg["ttl"] - 1
            # who has the message or is about to get it: a list on the wire ("seen", an extension field), ONE shared set object
            # between the nodes of a process, so that co-resident hops see what their siblings already targeted
            known = msg.get("_seen_set")
            if known is None:
                known = set(msg.get("seen") or ())
                known.add(msg["source"])
            known.add(self.addr)
            pending = [n for n in self._neighbors.direct_addrs() if n not in known]
            known.update(pending)
            fwd["_seen_set"] = known
            local = [n for n in pending if self._neighbors.is_local(n)] if Settings.GOSSIP_INLINE_FORWARD else []
            if local:  # next hop lives in this process: its handler is a function call, made on this thread
                self._gossiper.add_message(fwd, local, inline_ok=True)
            if len(local) < len(pending):
                lset = set(local)
                fwd["seen"] = list(known)
                self._gossiper.add_message(fwd, [n for n in pending if n not in lset])
        return None

    # ------------------------------------------------------------------ RPC adapters
    def _rpc_handshake(self, request: str, context) -> Optional[str]:
        return self.handshake(request)

    def _rpc_disconnect(self, request: str, context) -> None:
        self.disconnect(request)

    def _rpc_send(self, request: Dict[str, Any], context) -> Optional[str]:
        return self.send(request)


def _maybe_descriptor(weights: Any) -> Any:
    """Bytes that start with the descriptor magic are a pickled cross-process :class:`SlabRef`."""
    from myfyp_b200.parallel.ipc import decode_descriptor, is_descriptor

    if isinstance(weights, (bytes, bytearray)) and is_descriptor(weights):
        return decode_descriptor(bytes(weights))
    return weights
