"""Unix-domain-socket protocol: the control plane between the ranks of ONE box.

Same commands, flood, heartbeats and neighbour semantics as the other two transports (it reuses the gRPC protocol's
client / neighbour / servicer logic); only the pipe differs: length-prefixed pickled frames over ``AF_UNIX`` stream
sockets (``multiprocessing.connection``), one persistent connection per neighbour; weights and handshakes are
request -> response, control messages are one-way frames.  A Python gRPC
unary call costs 0.3-1 ms of interpreter time on each side; a frame on a unix socket costs tens of microseconds, which
is what a federation needs when every round is a few hundred small control messages and the model bytes travel over
NVLink anyway (weights messages carry slab descriptors, see :mod:`myfyp_b200.parallel.ipc`).

Not in the reference (it has gRPC and in-memory only).  Pickle over a local socket: use between processes you own.
"""

from __future__ import annotations

import os
import random
import threading
from multiprocessing.connection import Client as _ConnClient
from multiprocessing.connection import Listener
from typing import Any, List, Optional, Tuple

from myfyp_b200.communication.commands.command import Command
from myfyp_b200.communication.protocols.grpc.grpc_client import GrpcClient
from myfyp_b200.communication.protocols.grpc.grpc_communication_protocol import GrpcCommunicationProtocol
from myfyp_b200.communication.protocols.grpc.grpc_neighbors import GrpcNeighbors, LocalStub
from myfyp_b200.communication.protocols.grpc.grpc_server import GrpcServer, _local_lock, _local_servers, local_server
from myfyp_b200.management.logger import logger
from myfyp_b200.settings import Settings

SCHEME = "uds://"


def _path(addr: str) -> str:
    return addr[len(SCHEME) :] if addr.startswith(SCHEME) else addr


class UdsStub:
    """One persistent connection to a peer's listener; calls are request -> response, serialised per connection."""

    local = False

    def __init__(self, addr: str) -> None:
        self._conn = _ConnClient(_path(addr), family="AF_UNIX")
        self._lock = threading.Lock()

    def _call(self, kind: str, payload: Any, timeout: float) -> Any:
        with self._lock:
            self._conn.send((kind, payload))
            if not self._conn.poll(timeout):
                raise TimeoutError(f"no answer within {timeout} s")
            return self._conn.recv()

    def handshake(self, addr: str, timeout: float) -> Optional[str]:
        return self._call("h", addr, timeout)

    def disconnect(self, addr: str, timeout: float) -> None:
        self._call("d", addr, timeout)

    def send(self, msg: Any, timeout: float) -> Optional[str]:
        return self._call("s", msg, timeout)

    def send_oneway(self, msg: Any) -> None:
        """Control messages (votes, announcements, metrics, beats) whose sender ignores the peer's verdict travel as one-way
        frames: waiting for the (error-text-only) answer made every flood fan-out cost one round trip per neighbour.  Frames
        of one connection are handled in order, so later request/response calls still pair up; a dead peer shows up as a
        socket error here or through its missing heartbeats."""
        with self._lock:
            self._conn.send(("a", msg))
        return None

    def close(self) -> None:
        try:
            self._conn.close()
        except Exception:  # noqa: BLE001
            pass


class UdsNeighbors(GrpcNeighbors):
    def make_stub(self, addr: str) -> Tuple[Any, Any]:
        if Settings.GRPC_LOCAL_SHORTCUT and local_server(addr) is not None:
            return None, LocalStub(addr)
        stub = UdsStub(addr)
        return stub, stub  # the stub is its own closable "channel"


class UdsServer(GrpcServer):
    """The gRPC servicer logic (dedupe -> dispatch -> re-gossip) behind a unix-socket listener."""

    def start(self, wait: bool = False) -> None:
        path = _path(self.addr)
        try:
            os.unlink(path)
        except OSError:
            pass
        self._listener = Listener(path, family="AF_UNIX", backlog=64)
        self._started = True
        self._terminated.clear()
        with _local_lock:
            _local_servers[self.addr] = self
        threading.Thread(target=self._accept_loop, daemon=True, name=f"uds-accept-{self.addr}").start()
        logger.debug(self.addr, f"unix-socket server started at {self.addr}")
        if wait:
            self.wait_for_termination()

    def _accept_loop(self) -> None:
        while self._started:
            try:
                conn = self._listener.accept()
            except Exception:  # noqa: BLE001 - listener closed
                return
            threading.Thread(target=self._serve, args=(conn,), daemon=True, name=f"uds-conn-{self.addr}").start()

    def _serve(self, conn) -> None:
        try:
            while self._started:
                try:
                    kind, payload = conn.recv()
                except (EOFError, OSError):
                    return
                if kind == "a":  # one-way control frame: no answer is written
                    try:
                        err = self.send(payload)
                        if err:
                            logger.debug(self.addr, f"one-way frame refused: {err}")
                    except Exception as e:  # noqa: BLE001
                        logger.debug(self.addr, f"one-way frame failed: {e}")
                    continue
                try:
                    if kind == "s":
                        res = self.send(payload)
                    elif kind == "h":
                        res = self.handshake(payload)
                    elif kind == "d":
                        self.disconnect(payload)
                        res = None
                    else:
                        res = f"unknown frame kind {kind!r}"
                except Exception as e:  # noqa: BLE001
                    res = f"server error: {e}"
                try:
                    conn.send(res)
                except (OSError, ValueError):
                    return
        finally:
            try:
                conn.close()
            except Exception:  # noqa: BLE001
                pass

    def stop(self) -> None:
        with _local_lock:
            if _local_servers.get(self.addr) is self:
                del _local_servers[self.addr]
        self._started = False
        try:
            self._listener.close()
        except Exception:  # noqa: BLE001
            pass
        try:
            os.unlink(_path(self.addr))
        except OSError:
            pass
        self._terminated.set()


class UdsCommunicationProtocol(GrpcCommunicationProtocol):
    """Same servicer / client / neighbour logic as the gRPC transport, over a unix-socket pipe (and the same fabric test)."""

    def __init__(self, addr: Optional[str] = None, commands: Optional[List[Command]] = None) -> None:
        if not addr or addr == "127.0.0.1":
            addr = f"{SCHEME}/tmp/myfyp-b200-{os.getpid()}-{random.randint(0, 1 << 30)}.sock"
        elif not addr.startswith(SCHEME):
            addr = SCHEME + addr
        self._assemble(addr, UdsNeighbors(addr), GrpcClient, UdsServer, commands)
