"""Process-wide address -> in-memory server registry.

Reference: communication/protocols/memory/server_singleton.py:22-43.  Unlike the
reference, stopping one server removes *only that server*  - the reference
drops the whole registry, which breaks every other node in the process.
"""

from __future__ import annotations

import threading
from typing import Dict, Optional


class ServerRegistry:
    _servers: Dict[str, "object"] = {}
    _lock = threading.Lock()
    _changed = threading.Condition(_lock)

    @classmethod
    def register(cls, addr: str, server) -> None:
        with cls._changed:
            cls._servers[addr] = server
            cls._changed.notify_all()

    @classmethod
    def unregister(cls, addr: str, server=None) -> None:
        with cls._changed:
            if server is None or cls._servers.get(addr) is server:
                cls._servers.pop(addr, None)
            cls._changed.notify_all()

    @classmethod
    def get(cls, addr: str):
        srv = cls._servers.get(addr)
        if srv is None:
            raise KeyError(f"no in-memory server at {addr}")
        return srv

    @classmethod
    def lookup(cls, addr: str) -> Optional[object]:
        return cls._servers.get(addr)

    @classmethod
    def wait_until_gone(cls, addr: str) -> None:
        with cls._changed:
            cls._changed.wait_for(lambda: addr not in cls._servers)

    @classmethod
    def reset(cls) -> None:
        """Test helper: forget everything."""
        with cls._changed:
            cls._servers.clear()
            cls._changed.notify_all()


# name kept for scripts written against the reference
ServerSingleton = ServerRegistry
