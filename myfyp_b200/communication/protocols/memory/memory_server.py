"""In-memory server: handlers run synchronously on the *sender's* thread.

Reference: communication/protocols/memory/memory_server.py:40-222.
"""

from __future__ import annotations

import time
from typing import Any, Dict, List, Optional, Union

from myfyp_b200.communication.commands.command import Command
from myfyp_b200.communication.protocols.gossiper import Gossiper
from myfyp_b200.communication.protocols.memory.registry import ServerRegistry
from myfyp_b200.management.logger import logger
from myfyp_b200.settings import Settings


class InMemoryServer:
    def __init__(self, addr: str, gossiper: Gossiper, neighbors, commands: Optional[List[Command]] = None) -> None:
        self.addr = addr
        self.__commands: Dict[str, Command] = {c.get_name(): c for c in (commands or [])}
        self.__gossiper = gossiper
        self.__neighbors = neighbors
        self.__started = False

    # ------------------------------------------------------------------ lifecycle
    def start(self, wait: bool = False) -> None:
        ServerRegistry.register(self.addr, self)
        self.__started = True
        logger.debug(self.addr, f"InMemoryServer started at {self.addr}")
        if wait:
            self.wait_for_termination()

    def stop(self) -> None:
        ServerRegistry.unregister(self.addr, self)
        self.__started = False
        logger.debug(self.addr, f"InMemoryServer stopped at {self.addr}")

    def wait_for_termination(self) -> None:
        ServerRegistry.wait_until_gone(self.addr)

    def is_running(self) -> bool:
        return self.__started

    @property
    def gossiper(self) -> Gossiper:
        return self.__gossiper

    # ------------------------------------------------------------------ services
    def handshake(self, request: Dict[str, Any]) -> Dict[str, Any]:
        if not self.__started:
            return {"error": "server not running"}
        if self.__neighbors.add(request["addr"], non_direct=False, handshake_msg=False) or self.__neighbors.exists(request["addr"]):
            return {}
        return {"error": "Cannot add the node (duplicated or wrong direction)"}

    def disconnect(self, request: Dict[str, Any]) -> None:
        self.__neighbors.remove(request["addr"], disconnect_msg=False)

    def send_message(self, request: Dict[str, Any]) -> Dict[str, Any]:
        if not self.__started:
            return {"error": "server not running"}
        if not self.__gossiper.check_and_set_processed(request["hash"]):
            return {}
        cmd = request["cmd"]
        if cmd != "beat":
            logger.log_bytes("msgs/" + cmd, 1)  # control-plane message census (benchmarks report it per round)
        if cmd != "beat" and request["source"] != self.addr:
            # any message proves its origin is alive: learn about it now rather than at its next heartbeat (a node that
            # just (re)joined would otherwise be missing from some nodes' vote bookkeeping -> diverging train sets)
            self.__neighbors.refresh_or_add(request["source"], time.time())
        if logger.is_debug() and (cmd != "beat" or not Settings.EXCLUDE_BEAT_LOGS):
            logger.debug(self.addr, f"📫 {cmd.upper()} received from {request['source']} (ttl={request['ttl']})")
        # re-gossip first (same order as the reference, memory_server.py:155-165)
        if request["ttl"] > 0:
            fwd = dict(request)
            fwd["ttl"] = request["ttl"] - 1
            # Flood bookkeeping inside one process: ONE set object travels with the message through every hop ("who has it or is
            # about to get it"), so a hop sees what its siblings have already targeted - the per-path lists of the wire protocols
            # (gRPC / UDS: "seen") let ~10 redundant probes through per delivery at 64 nodes with degree 20.
            known = request.get("_seen_set")
            if known is None:
                known = set(request.get("seen") or ())
                known.add(request["source"])
            known.add(self.addr)
            pending = [n for n in self.__neighbors.direct_addrs() if n not in known]
            known.update(pending)
            fwd["_seen_set"] = known
            self.__gossiper.add_message(fwd, pending, inline_ok=True)
        handler = self.__commands.get(cmd)
        if handler is None:
            logger.error(self.addr, f"Unknown command: {cmd} from {request['source']}")
            return {"error": f"Unknown command: {cmd}"}
        try:
            handler.execute(request["source"], request["round"], *request["args"])
        except Exception as e:  # noqa: BLE001
            text = f"Error while processing command: {cmd} {request['args']}: {e}"
            logger.error(self.addr, text)
            return {"error": text}
        return {}

    def send_weights(self, request: Dict[str, Any]) -> Dict[str, Any]:
        if not self.__started:
            return {"error": "server not running"}
        handler = self.__commands.get(request["cmd"])
        logger.log_bytes("msgs/" + request["cmd"], 1)
        if handler is None:
            logger.error(self.addr, f"Unknown command: {request['cmd']} from {request['source']}")
            return {"error": f"Unknown command: {request['cmd']}"}
        try:
            handler.execute(
                request["source"], request["round"], weights=request["weights"], contributors=request["contributors"], num_samples=request["weight"]
            )
        except Exception as e:  # noqa: BLE001
            text = f"Error while processing model: {request['cmd']}: {e}"
            logger.error(self.addr, text)
            return {"error": text}
        return {}

    def add_command(self, cmds: Union[Command, List[Command]]) -> None:
        for c in cmds if isinstance(cmds, list) else [cmds]:
            self.__commands[c.get_name()] = c
