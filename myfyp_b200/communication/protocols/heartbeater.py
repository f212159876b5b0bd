"""Heartbeater (reference: communication/protocols/heartbeater.py:33-113).

Every ``HEARTBEAT_PERIOD`` broadcast ``beat [time]`` and evict
neighbours silent for more than ``HEARTBEAT_TIMEOUT``.  ``beat()`` refreshes (or adds,
as non-direct) the neighbour a flooded heartbeat came from.  The very first tick only
sends (peers need one period to answer before they can be judged).
"""

from __future__ import annotations

import threading
import time


from myfyp_b200.communication.protocols.client import Client
from myfyp_b200.communication.protocols.neighbors import Neighbors
from myfyp_b200.management.logger import logger
from myfyp_b200.settings import Settings

heartbeater_cmd_name = "beat"


class Heartbeater(threading.Thread):
    def __init__(self, self_addr: str, neighbors: Neighbors, client: Client) -> None:
        super().__init__(daemon=True, name=f"heartbeater-thread-{self_addr}")
        self.__self_addr = self_addr
        self.__neighbors = neighbors
        self.__client = client
        self.__terminate = threading.Event()

    def stop(self) -> None:
        self.__terminate.set()

    def beat(self, nei: str, time: float) -> None:
        if nei == self.__self_addr:
            return
        self.__neighbors.refresh_or_add(nei, time)

    def run(self) -> None:
        toggle = False
        while not self.__terminate.is_set():
            t0 = time.monotonic()
            now = time.time()
            if toggle:
                timeout = Settings.HEARTBEAT_TIMEOUT
                for nei, entry in self.__neighbors.get_all().items():
                    if now - entry[2] > timeout:
                        logger.info(self.__self_addr, f"Heartbeat timeout for {nei} ({now - entry[2]:.2f}). Removing...")
                        direct = bool(entry[1])
                        self.__neighbors.remove(nei)
                        if direct:
                            self.__neighbors.report_lost(nei)
            toggle = True  # first tick only sends; every later tick also evicts
            try:
                self.__client.broadcast(self.__client.build_message(heartbeater_cmd_name, args=[str(time.time())]))
            except Exception as e:  # noqa: BLE001
                logger.debug(self.__self_addr, f"heartbeat broadcast failed: {e}")
            left = Settings.HEARTBEAT_PERIOD - (time.monotonic() - t0)
            self.__terminate.wait(max(left, 0.01))
