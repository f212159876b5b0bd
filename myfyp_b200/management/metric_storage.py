"""Thread-safe metric stores.

Shapes match the reference so user code reading the logs keeps working
(reference: p2pfl/management/metric_storage.py:30-251):

* local  : ``exp -> round -> node -> metric -> [(step, value), ...]``
* global : ``exp -> node -> metric -> [(round, value), ...]`` (one value per round)
"""

from __future__ import annotations

import copy
import threading
from typing import Dict, List, Tuple, Union

MetricsType = Dict[str, List[Tuple[int, float]]]
NodeLogsType = Dict[str, MetricsType]
RoundLogsType = Dict[int, NodeLogsType]
LocalLogsType = Dict[str, RoundLogsType]
GlobalLogsType = Dict[str, NodeLogsType]


class LocalMetricStorage:
    """Per-training-step metrics."""

    def __init__(self, disable_locks: bool = False) -> None:
        self._logs: LocalLogsType = {}
        self._lock = threading.Lock()

    def add_log(self, exp_name: str, round: int, metric: str, node: str, val: Union[int, float], step: int) -> None:
        with self._lock:
            series = self._logs.setdefault(exp_name, {}).setdefault(round, {}).setdefault(node, {}).setdefault(metric, [])
            series.append((step, val))

    def add_series(self, exp_name: str, round: int, metric: str, node: str, values, first_step: int) -> None:
        """``values[i]`` at step ``first_step + i`` - one lock round trip for a whole epoch of per-step values."""
        with self._lock:
            series = self._logs.setdefault(exp_name, {}).setdefault(round, {}).setdefault(node, {}).setdefault(metric, [])
            series.extend((first_step + i, v) for i, v in enumerate(values))

    def get_all_logs(self) -> LocalLogsType:
        with self._lock:
            return copy.deepcopy(self._logs)

    def get_experiment_logs(self, exp: str) -> RoundLogsType:
        with self._lock:
            return copy.deepcopy(self._logs[exp])

    def get_experiment_round_logs(self, exp: str, round: int) -> NodeLogsType:
        with self._lock:
            return copy.deepcopy(self._logs[exp][round])

    def get_experiment_round_node_logs(self, exp: str, round: int, node: str) -> MetricsType:
        with self._lock:
            return copy.deepcopy(self._logs[exp][round][node])

    def clear(self) -> None:
        with self._lock:
            self._logs = {}


class GlobalMetricStorage:
    """Per-round metrics (evaluation results, own and received from peers)."""

    def __init__(self, disable_locks: bool = False) -> None:
        self._logs: GlobalLogsType = {}
        self._lock = threading.Lock()

    def add_log(self, exp_name: str, round: int, metric: str, node: str, val: Union[int, float]) -> None:
        with self._lock:
            series = self._logs.setdefault(exp_name, {}).setdefault(node, {}).setdefault(metric, [])
            # a round holds one value: keep the first one logged for it (rounds arrive in order: looking at the tail is enough
            # in all but pathological cases, which still fall back to the full scan)
            if not series or (series[-1][0] != round and (series[-1][0] < round or round not in [r for r, _ in series])):
                series.append((round, val))

    def get_all_logs(self) -> GlobalLogsType:
        with self._lock:
            return copy.deepcopy(self._logs)

    def get_experiment_logs(self, exp: str) -> NodeLogsType:
        with self._lock:
            return copy.deepcopy(self._logs[exp])

    def get_experiment_node_logs(self, exp: str, node: str) -> MetricsType:
        with self._lock:
            return copy.deepcopy(self._logs[exp][node])

    def clear(self) -> None:
        with self._lock:
            self._logs = {}
