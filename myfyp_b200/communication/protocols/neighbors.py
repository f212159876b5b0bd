"""Neighbour table (reference: communication/protocols/neighbors.py:27-167).

``neis[addr] = (channel, stub_or_server, last_beat_time)``; a *direct* neighbour is
one whose second field is truthy.  Listeners registered with ``on_change`` are called
after every add/remove so waiting stages wake up without polling.
"""

from __future__ import annotations

import threading
from typing import Any, Callable, Dict, List

from myfyp_b200.management.logger import logger


class Neighbors:
    def __init__(self, self_addr: str) -> None:
        self.self_addr = self_addr
        self.neis: Dict[str, Any] = {}
        self.neis_lock = threading.RLock()
        self._listeners: List[Callable[[], None]] = []
        self._direct: Any = None  # cached tuple of the direct neighbours' addresses (membership changes only in add / remove)
        self._lost_listeners: List[Callable[[str], None]] = []
        self._tombstones: Dict[str, float] = {}  # addr -> monotonic time until which stray messages do not re-add it

    # -- to be provided by the transport ------------------------------------------
    def connect(self, addr: str, *args, **kwargs) -> Any:
        raise NotImplementedError

    def disconnect(self, addr: str, *args, **kwargs) -> None:
        raise NotImplementedError

    def refresh_or_add(self, addr: str, time: float) -> None:
        raise NotImplementedError

    # -- generic ---------------------------------------------------------------
    def on_change(self, fn: Callable[[], None]) -> None:
        self._listeners.append(fn)

    def _changed(self) -> None:
        self._direct = None
        for fn in self._listeners:
            try:
                fn()
            except Exception:
                pass

    # -- failure dissemination --------------------------------------------------
    def on_lost(self, fn: Callable[[str], None]) -> None:
        """``fn(addr)`` is called when THIS node found a direct neighbour dead (a send to it failed, or its beats stopped)."""
        self._lost_listeners.append(fn)

    def report_lost(self, addr: str) -> None:
        for fn in self._lost_listeners:
            try:
                fn(addr)
            except Exception:  # noqa: BLE001
                pass

    def bury(self, addr: str, seconds: float) -> bool:
        """Somebody else found ``addr`` dead (``node_down``): forget it if we only know it through the flood, and do not let its
        messages that are still travelling re-add it for ``seconds``.  A direct neighbour is left to our own link (the next beat we
        send it fails or does not); a node declared dead by mistake comes back with its first beat after the grace period."""
        import time as _t

        with self.neis_lock:
            cur = self.neis.get(addr)
            if cur is None or cur[1]:
                return False
            self._tombstones[addr] = _t.monotonic() + seconds
        self.remove(addr, disconnect_msg=False)
        return True

    def buried(self, addr: str) -> bool:
        t = self._tombstones.get(addr)
        if t is None:
            return False
        import time as _t

        if _t.monotonic() >= t:
            self._tombstones.pop(addr, None)
            return False
        return True

    def add(self, addr: str, *args, **kwargs) -> bool:
        if addr == self.self_addr:
            logger.info(self.self_addr, "❌ Cannot add itself")
            return False
        if kwargs.get("non_direct") and self.buried(addr):
            return False
        with self.neis_lock:
            cur = self.neis.get(addr)
            # a node only known through the heartbeat flood (non-direct) may be upgraded to a direct neighbour by an explicit
            # connect; the reference refuses that as a duplicate (neighbors.py:85-88), which makes a topology impossible to densify
            if cur is not None and (cur[1] or kwargs.get("non_direct")):
                logger.debug(self.self_addr, f"❌ Cannot add duplicates. {addr} already exists.")
                return False
        # connect outside the lock: the handshake calls back into the peer, which may
        # concurrently be connecting to us (two locks, opposite order => deadlock)
        try:
            entry = self.connect(addr, *args, **kwargs)
        except Exception as e:  # noqa: BLE001
            logger.info(self.self_addr, f"❌ Cannot add {addr}: {e}")
            return False
        with self.neis_lock:
            cur = self.neis.get(addr)
            if cur is None or (not cur[1] and entry[1]):
                self.neis[addr] = entry
                self._direct = None
        self._changed()
        return True

    def remove(self, addr: str, *args, **kwargs) -> None:
        with self.neis_lock:
            present = addr in self.neis
            entry = self.neis.pop(addr, None)
            self._direct = None
        if present:
            try:
                self.disconnect(addr, *args, entry=entry, **kwargs)
            except Exception as e:  # noqa: BLE001
                logger.debug(self.self_addr, f"Error while disconnecting from {addr}: {e}")
            self._changed()

    def get(self, addr: str) -> Any:
        return self.neis[addr]

    def is_direct(self, addr: str) -> bool:
        e = self.neis.get(addr)
        return bool(e is not None and e[1])

    def is_local(self, addr: str) -> bool:
        """True when ``addr`` is a direct neighbour living in this process (its server is called, not messaged)."""
        e = self.neis.get(addr)
        return bool(e is not None and e[1] is not None and getattr(e[1], "local", False))

    def direct_addrs(self) -> tuple:
        """Addresses of the direct neighbours - the flood's fan-out, asked for once per forwarded message."""
        d = self._direct
        if d is None:
            with self.neis_lock:
                d = tuple(k for k, v in self.neis.items() if v[1])
                self._direct = d
        return d

    def get_all(self, only_direct: bool = False) -> Dict[str, Any]:
        with self.neis_lock:
            neis = dict(self.neis)
        if only_direct:
            return {k: v for k, v in neis.items() if v[1]}
        return neis

    def exists(self, addr: str) -> bool:
        return addr in self.neis

    def clear_neighbors(self) -> None:
        for addr in list(self.neis.keys()):
            self.remove(addr)
