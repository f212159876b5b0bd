"""Resource monitor thread (reference: p2pfl/management/node_monitor.py:31-86).

Every ``Settings.RESOURCE_MONITOR_PERIOD`` seconds reports cpu %, ram %, network
in/out MB/s through a callback; on a GPU box it also reports per-GPU utilisation,
memory and power from NVML (B200 addition).
"""

from __future__ import annotations

import datetime
import threading
import time
from typing import Callable

from myfyp_b200.settings import Settings


class NodeMonitor(threading.Thread):
    def __init__(self, node_addr: str, report_fn: Callable[[str, str, float, datetime.datetime], None]) -> None:
        super().__init__(daemon=True, name=f"resource-monitor-thread-{node_addr}")
        self.node_addr = node_addr
        self.report_fn = report_fn
        self.period = Settings.RESOURCE_MONITOR_PERIOD
        self._stop_evt = threading.Event()

    def stop(self) -> None:
        self._stop_evt.set()

    def sample(self) -> dict:
        import psutil

        out = {"cpu": psutil.cpu_percent(), "ram": psutil.virtual_memory().percent}
        net = psutil.net_io_counters()
        out["_net_in"], out["_net_out"] = net.bytes_recv, net.bytes_sent
        out.update(gpu_metrics())
        return out

    def run(self) -> None:
        prev = self.sample()
        while not self._stop_evt.wait(self.period):
            cur = self.sample()
            now = datetime.datetime.now()
            for k, v in cur.items():
                if not k.startswith("_"):
                    self.report_fn(self.node_addr, k, float(v), now)
            dt = max(self.period, 1e-6)
            self.report_fn(self.node_addr, "net_in", (cur["_net_in"] - prev["_net_in"]) / dt / 2**20, now)
            self.report_fn(self.node_addr, "net_out", (cur["_net_out"] - prev["_net_out"]) / dt / 2**20, now)
            prev = cur


def gpu_metrics() -> dict:
    """NVML snapshot ``{gpu<i>_util, gpu<i>_mem_gb, gpu<i>_power_w, gpu<i>_sm_mhz}``; {} without a GPU."""
    try:
        import pynvml

        pynvml.nvmlInit()
        out = {}
        for i in range(pynvml.nvmlDeviceGetCount()):
            h = pynvml.nvmlDeviceGetHandleByIndex(i)
            out[f"gpu{i}_util"] = pynvml.nvmlDeviceGetUtilizationRates(h).gpu
            out[f"gpu{i}_mem_gb"] = pynvml.nvmlDeviceGetMemoryInfo(h).used / 2**30
            out[f"gpu{i}_power_w"] = pynvml.nvmlDeviceGetPowerUsage(h) / 1000.0
            out[f"gpu{i}_sm_mhz"] = pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)
        return out
    except Exception:
        return {}
