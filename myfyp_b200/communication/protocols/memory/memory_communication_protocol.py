"""In-memory transport: every node of the process registers its server in ``ServerRegistry`` and a "send" is a method call
on the peer's server object - nothing is serialised, weights travel as device descriptors.
(Reference: communication/protocols/memory/memory_communication_protocol.py:64-268; the facade itself is ``ComposedProtocol``.)"""

from __future__ import annotations

import random
from typing import List, Optional

from myfyp_b200.communication.commands.command import Command
from myfyp_b200.communication.protocols.composed_protocol import ComposedProtocol
from myfyp_b200.communication.protocols.memory.memory_client import InMemoryClient
from myfyp_b200.communication.protocols.memory.memory_neighbors import InMemoryNeighbors
from myfyp_b200.communication.protocols.memory.memory_server import InMemoryServer
from myfyp_b200.communication.protocols.memory.registry import ServerRegistry


class InMemoryCommunicationProtocol(ComposedProtocol):
    def __init__(self, addr: Optional[str] = None, commands: Optional[List[Command]] = None) -> None:
        # Node() passes its default "127.0.0.1": for an in-process protocol that means "pick a name"
        name = addr if addr and addr != "127.0.0.1" else f"node-{random.randint(0, 99999999)}"
        self._assemble(name, InMemoryNeighbors(name), InMemoryClient, InMemoryServer, commands)

    def fabric_reachable(self, nei: str) -> bool:
        """Same process => the peer's slab is dereferenceable (same GPU or via P2P)."""
        return ServerRegistry.lookup(nei) is not None
