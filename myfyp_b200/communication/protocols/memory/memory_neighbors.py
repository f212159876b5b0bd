"""In-memory neighbour table (reference: communication/protocols/memory/memory_neighbors.py:30-109)."""

from __future__ import annotations

import time as _time
from typing import Any, Optional, Tuple

from myfyp_b200.communication.protocols.memory.registry import ServerRegistry
from myfyp_b200.communication.protocols.neighbors import Neighbors
from myfyp_b200.management.logger import logger


class InMemoryNeighbors(Neighbors):
    def refresh_or_add(self, addr: str, time: float) -> None:
        with self.neis_lock:
            cur = self.neis.get(addr)
            if cur is not None:
                self.neis[addr] = (cur[0], cur[1], time)
                return
        self.add(addr, non_direct=True)

    def connect(self, addr: str, non_direct: bool = False, handshake_msg: bool = True) -> Tuple[None, Optional[Any], float]:
        if non_direct:
            logger.debug(self.self_addr, f"🔍 Found node {addr}")
            return (None, None, _time.time())
        logger.debug(self.self_addr, f"🤝 Adding {addr}")
        server = ServerRegistry.get(addr)  # KeyError => add() returns False
        if handshake_msg:
            res = server.handshake({"addr": self.self_addr})
            if res.get("error"):
                raise Exception(f"Cannot add a neighbor: {res['error']}")
        return (None, server, _time.time())

    def disconnect(self, addr: str, disconnect_msg: bool = True, entry: Any = None) -> None:
        if entry is None:
            entry = self.neis.get(addr)
        if entry is None:
            return
        server = entry[1]
        if disconnect_msg and server is not None:
            server.disconnect({"addr": self.self_addr})
