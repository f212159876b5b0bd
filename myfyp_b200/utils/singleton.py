"""``SingletonMeta`` (reference: utils/singleton.py:22-32): one instance per class, thread-safe."""

from __future__ import annotations

import threading
from typing import Any, Dict


class SingletonMeta(type):
    _instances: Dict[type, Any] = {}
    _lock = threading.Lock()

    def __call__(cls, *args, **kwargs):
        if cls not in SingletonMeta._instances:
            with SingletonMeta._lock:
                if cls not in SingletonMeta._instances:
                    SingletonMeta._instances[cls] = super().__call__(*args, **kwargs)
        return SingletonMeta._instances[cls]
