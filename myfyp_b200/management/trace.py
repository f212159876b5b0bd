"""Opt-in event timeline of the control plane (``B2_TRACE=1``): ``(monotonic seconds, node, event)`` tuples in memory.

For finding out where a federated round waits: which message was sent when, how long a handler ran, when a node's
aggregation completed.  ``CLOCK_MONOTONIC`` is shared by the processes of one box, so the per-rank dumps line up.
Not in the reference (its only timing facility is the log file).
"""

from __future__ import annotations

import json
import os
import threading
import time
from typing import List, Tuple

ENABLED = os.environ.get("B2_TRACE", "0") == "1"
_events: List[Tuple[float, str, str]] = []
_lock = threading.Lock()


def trace(node: str, event: str) -> None:
    if ENABLED:
        with _lock:
            _events.append((time.monotonic(), node, event))


def clear() -> None:
    with _lock:
        _events.clear()


def dump(path: str) -> None:
    with _lock:
        ev = list(_events)
    with open(path, "w") as f:
        json.dump(ev, f)
