"""The ``logger`` singleton.

Public surface follows the reference's logger (p2pfl/management/logger/logger.py:96-454):
``info/debug/warning/error/critical(node, msg)``, ``log_metric``,
``get_local_logs/get_global_logs``, ``register_node/unregister_node``,
``experiment_started/finished``, ``round_started/finished``, ``set_level/get_level``,
``connect_web``, ``cleanup``.

The reference composes a stack of decorator objects at import time
(file -> async queue -> web -> singleton).  Here it is one class with pluggable
``logging`` handlers: a coloured stream handler by default, an optional rotating
file handler (``Settings.FILE_LOGGING``), an optional web push (``connect_web``).
B200 additions: per-round device timings and bytes-moved counters
(``log_device_time`` / ``log_bytes``) and GPU system metrics from NVML.
"""

from __future__ import annotations

import atexit
import datetime
import logging
import os
import threading
from logging.handlers import RotatingFileHandler
from typing import Any, Dict, List, Optional, Union

from myfyp_b200.experiment import Experiment
from myfyp_b200.management.metric_storage import GlobalLogsType, GlobalMetricStorage, LocalLogsType, LocalMetricStorage
from myfyp_b200.settings import Settings

_GRAY, _YELLOW, _RED, _BLUE, _CYAN, _RESET = "\033[90m", "\033[93m", "\033[91m", "\033[94m", "\033[96m", "\033[0m"


class NodeNotRegistered(Exception):
    """Operation on a node the logger does not know."""


class _ColourFormatter(logging.Formatter):
    _COLOURS = {"DEBUG": _BLUE, "INFO": _CYAN, "WARNING": _YELLOW, "ERROR": _RED, "CRITICAL": _RED}

    def format(self, record: logging.LogRecord) -> str:
        c = self._COLOURS.get(record.levelname, _RESET)
        ts = datetime.datetime.fromtimestamp(record.created).strftime("%Y-%m-%d %H:%M:%S")
        node = getattr(record, "node", "-")
        return f"{_GRAY}[ {_YELLOW}{ts}{_GRAY} | {_CYAN}{node}{_GRAY} | {c}{record.levelname}{_GRAY} ]{_RESET} {record.getMessage()}"


class P2PFLogger:
    """Process-wide logger + metric stores."""

    def __init__(self) -> None:
        self._logger = logging.getLogger("myfyp_b200")
        self._logger.propagate = False
        self._logger.setLevel(logging.getLevelName(Settings.LOG_LEVEL))
        if not self._logger.handlers:
            h = logging.StreamHandler()
            h.setFormatter(_ColourFormatter())
            self._logger.addHandler(h)
        self._file_handler: Optional[logging.Handler] = None
        self._nodes: Dict[str, Dict[str, Any]] = {}
        self._nodes_lock = threading.Lock()
        self.local_metrics = LocalMetricStorage()
        self.global_metrics = GlobalMetricStorage()
        # device-side accounting (B200 addition)
        self._device_times: Dict[str, List[float]] = {}
        self._bytes: Dict[str, int] = {}
        self._acct_lock = threading.Lock()
        self._web = None
        self._monitors: Dict[str, Any] = {}
        if Settings.FILE_LOGGING:
            self.enable_file_logging()
        atexit.register(self.cleanup)

    # ------------------------------------------------------------------ handlers
    def enable_file_logging(self, log_dir: Optional[str] = None) -> None:
        """Rotating ``<LOG_DIR>/myfyp_b200.log`` (1 MB x 3), like the reference's FileLogger."""
        if self._file_handler is not None:
            return
        log_dir = log_dir or Settings.LOG_DIR
        os.makedirs(log_dir, exist_ok=True)
        h = RotatingFileHandler(os.path.join(log_dir, "myfyp_b200.log"), maxBytes=1_000_000, backupCount=3)
        h.setFormatter(logging.Formatter("[ %(asctime)s | %(node)s | %(levelname)s ]: %(message)s", defaults={"node": "-"}))
        self._logger.addHandler(h)
        self._file_handler = h

    def add_handler(self, handler: logging.Handler) -> None:
        self._logger.addHandler(handler)

    def enable_async(self) -> None:
        """Hand records to a background thread through a queue (the reference's AsyncLogger decorator,
        management/logger/decorators/async_logger.py:29-58): formatting and I/O leave the learning threads."""
        import queue
        from logging.handlers import QueueHandler, QueueListener

        if getattr(self, "_listener", None) is not None:
            return
        q: "queue.Queue" = queue.Queue(-1)
        handlers = list(self._logger.handlers)
        for h in handlers:
            self._logger.removeHandler(h)
        self._listener = QueueListener(q, *handlers, respect_handler_level=True)
        self._listener.start()
        self._logger.addHandler(QueueHandler(q))
        atexit.register(self._listener.stop)

    def connect_web(self, url: str, key: str) -> None:
        """Push logs/metrics to a P2PFL-web-services-compatible REST endpoint."""
        from myfyp_b200.management.web_services import P2pflWebServices, WebLogHandler

        self._web = P2pflWebServices(url, key)
        self.add_handler(WebLogHandler(self._web))

    def cleanup(self) -> None:
        for m in list(self._monitors.values()):
            try:
                m.stop()
            except Exception:
                pass
        self._monitors.clear()
        for h in list(self._logger.handlers):
            try:
                h.flush()
            except Exception:
                pass

    # ------------------------------------------------------------------ levels
    def set_level(self, level: Union[int, str]) -> None:
        if isinstance(level, str):
            level = logging.getLevelName(level.upper())
            if not isinstance(level, int):
                raise ValueError(f"Invalid level: {level}")
        self._logger.setLevel(level)

    def get_level(self) -> int:
        return self._logger.getEffectiveLevel()

    def get_level_name(self, lvl: int) -> str:
        return logging.getLevelName(lvl)

    # ------------------------------------------------------------------ text logs
    def log(self, level: int, node: str, message: str) -> None:
        if self._logger.isEnabledFor(level):
            self._logger.log(level, message, extra={"node": node})

    def info(self, node: str, message: str) -> None:
        self.log(logging.INFO, node, message)

    def is_debug(self) -> bool:
        """Hot paths ask before formatting a debug line."""
        return self._logger.isEnabledFor(logging.DEBUG)

    def debug(self, node: str, message: str) -> None:
        self.log(logging.DEBUG, node, message)

    def warning(self, node: str, message: str) -> None:
        self.log(logging.WARNING, node, message)

    def error(self, node: str, message: str) -> None:
        self.log(logging.ERROR, node, message)

    def critical(self, node: str, message: str) -> None:
        self.log(logging.CRITICAL, node, message)

    # ------------------------------------------------------------------ metrics
    def log_metric(self, addr: str, metric: str, value: float, round: Optional[int] = None, step: Optional[int] = None) -> None:
        """Store a metric.  ``step is None`` => global (per round) store, else local store.

        Like the reference (logger.py:286-308) a metric for a node with no
        registered experiment is dropped silently and the node's *own* experiment round
        is used, whatever ``round`` says.
        """
        try:
            experiment = self._nodes[addr]["Experiment"]
        except KeyError:
            return
        if experiment is None or experiment.round is None or experiment.exp_name is None:
            return
        if step is None:
            self.global_metrics.add_log(experiment.exp_name, experiment.round, metric, addr, value)
        else:
            self.local_metrics.add_log(experiment.exp_name, experiment.round, metric, addr, value, step)
        if self._web is not None:
            try:
                if step is None:
                    self._web.send_global_metric(experiment.exp_name, experiment.round, metric, addr, value)
                else:
                    self._web.send_local_metric(experiment.exp_name, experiment.round, metric, addr, value, step)
            except Exception:
                pass

    def log_metric_series(self, addr: str, metric: str, values, first_step: int) -> None:
        """A learner's per-step values of one fit (98 losses per epoch in the benchmark) in one call: same stores and the same
        drop rule as :meth:`log_metric`, one lock acquisition instead of one per value."""
        try:
            experiment = self._nodes[addr]["Experiment"]
        except KeyError:
            return
        if experiment is None or experiment.round is None or experiment.exp_name is None or not len(values):
            return
        self.local_metrics.add_series(experiment.exp_name, experiment.round, metric, addr, values, first_step)
        if self._web is not None:
            for i, v in enumerate(values):
                try:
                    self._web.send_local_metric(experiment.exp_name, experiment.round, metric, addr, v, first_step + i)
                except Exception:
                    pass

    def log_system_metric(self, node: str, metric: str, value: float, time: datetime.datetime) -> None:
        if self._web is not None:
            try:
                self._web.send_system_metric(node, metric, value, time)
            except Exception:
                pass

    def get_local_logs(self) -> LocalLogsType:
        return self.local_metrics.get_all_logs()

    def get_global_logs(self) -> GlobalLogsType:
        return self.global_metrics.get_all_logs()

    # ------------------------------------------------------------------ device accounting (B200 addition)
    def log_device_time(self, key: str, ms: float) -> None:
        with self._acct_lock:
            self._device_times.setdefault(key, []).append(ms)

    def log_bytes(self, key: str, nbytes: int) -> None:
        with self._acct_lock:
            self._bytes[key] = self._bytes.get(key, 0) + int(nbytes)

    def get_device_times(self) -> Dict[str, List[float]]:
        with self._acct_lock:
            return {k: list(v) for k, v in self._device_times.items()}

    def get_bytes(self) -> Dict[str, int]:
        with self._acct_lock:
            return dict(self._bytes)

    # ------------------------------------------------------------------ node registry
    def register_node(self, node: str, simulation: bool = False) -> None:
        with self._nodes_lock:
            if node in self._nodes:
                raise Exception(f"Node {node} already registered.")
            self._nodes[node] = {"Experiment": None, "simulation": simulation}
        if self._web is not None:
            try:
                self._web.register_node(node, simulation)
                from myfyp_b200.management.node_monitor import NodeMonitor

                mon = NodeMonitor(node, self.log_system_metric)
                mon.start()
                self._monitors[node] = mon
            except Exception:
                pass

    def unregister_node(self, node: str) -> None:
        with self._nodes_lock:
            if node not in self._nodes:
                raise NodeNotRegistered(f"Node {node} not registered.")
            del self._nodes[node]
        mon = self._monitors.pop(node, None)
        if mon is not None:
            mon.stop()

    def get_nodes(self) -> Dict[str, Dict[Any, Any]]:
        return self._nodes

    # ------------------------------------------------------------------ lifecycle hooks
    def experiment_started(self, node: str, experiment: Optional[Experiment]) -> None:
        self.debug(node, "Experiment started.")
        if node in self._nodes:
            self._nodes[node]["Experiment"] = experiment

    def experiment_finished(self, node: str) -> None:
        self.debug(node, "Experiment finished.")

    def round_started(self, node: str, experiment: Optional[Experiment]) -> None:
        self.debug(node, f"Round {experiment.round if experiment else None} started.")
        if node in self._nodes:
            self._nodes[node]["Experiment"] = experiment

    def round_finished(self, node: str) -> None:
        self.debug(node, "Round finished.")

    # ------------------------------------------------------------------ test helper
    def reset_metrics(self) -> None:
        self.local_metrics.clear()
        self.global_metrics.clear()
        with self._acct_lock:
            self._device_times.clear()
            self._bytes.clear()


logger = P2PFLogger()
