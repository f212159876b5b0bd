"""Experiment helpers (reference: p2pfl/utils/utils.py:39-145)."""

from __future__ import annotations

import time
from typing import List

import numpy as np

from myfyp_b200.management.logger import logger
from myfyp_b200.settings import Settings


def set_test_settings() -> None:
    """Short timeouts for tests (same values as the reference's preset, utils.py:39-57)."""
    Settings.GRPC_TIMEOUT = 0.5
    Settings.HEARTBEAT_PERIOD = 0.5
    Settings.HEARTBEAT_TIMEOUT = 2
    Settings.GOSSIP_PERIOD = 0
    Settings.TTL = 10
    Settings.GOSSIP_MESSAGES_PER_PERIOD = 100
    Settings.AMOUNT_LAST_MESSAGES_SAVED = 100
    Settings.GOSSIP_MODELS_PERIOD = 1
    Settings.GOSSIP_MODELS_PER_ROUND = 4
    Settings.GOSSIP_EXIT_ON_X_EQUAL_ROUNDS = 4
    Settings.TRAIN_SET_SIZE = 4
    Settings.VOTE_TIMEOUT = 60
    Settings.AGGREGATION_TIMEOUT = 60
    Settings.WAIT_HEARTBEATS_CONVERGENCE = 0.2 * Settings.HEARTBEAT_TIMEOUT
    Settings.LOG_LEVEL = "DEBUG"
    Settings.EXCLUDE_BEAT_LOGS = True
    Settings.USE_SSL = False
    logger.set_level("DEBUG")


def set_standalone_settings() -> None:
    """Many nodes in one process (reference: examples/mnist.py:43-70)."""
    Settings.GRPC_TIMEOUT = 10
    Settings.HEARTBEAT_PERIOD = 1
    Settings.HEARTBEAT_TIMEOUT = 10
    Settings.GOSSIP_PERIOD = 0
    Settings.TTL = 10
    Settings.GOSSIP_MESSAGES_PER_PERIOD = 9999999999
    Settings.AMOUNT_LAST_MESSAGES_SAVED = 10000
    Settings.GOSSIP_MODELS_PERIOD = 1
    Settings.GOSSIP_MODELS_PER_ROUND = 4
    Settings.GOSSIP_EXIT_ON_X_EQUAL_ROUNDS = 10
    Settings.TRAIN_SET_SIZE = 4
    Settings.VOTE_TIMEOUT = 60
    Settings.AGGREGATION_TIMEOUT = 60
    Settings.WAIT_HEARTBEATS_CONVERGENCE = 0.2 * Settings.HEARTBEAT_TIMEOUT
    Settings.LOG_LEVEL = "INFO"
    Settings.EXCLUDE_BEAT_LOGS = True
    logger.set_level("INFO")


def set_fast_settings(train_set_size: int = 4) -> None:
    """Event-driven preset for benchmarks on one box: control-plane periods shrink to what the wake-ups need."""
    Settings.GRPC_TIMEOUT = 10
    Settings.HEARTBEAT_PERIOD = 1
    Settings.HEARTBEAT_TIMEOUT = 30
    Settings.GOSSIP_PERIOD = 0
    Settings.TTL = 10
    Settings.GOSSIP_MESSAGES_PER_PERIOD = 10_000_000
    Settings.AMOUNT_LAST_MESSAGES_SAVED = 10000
    Settings.GOSSIP_MODELS_PERIOD = 0.05
    Settings.GOSSIP_MODELS_PER_ROUND = 64
    Settings.GOSSIP_EXIT_ON_X_EQUAL_ROUNDS = 200
    Settings.TRAIN_SET_SIZE = train_set_size
    Settings.VOTE_TIMEOUT = 60
    Settings.AGGREGATION_TIMEOUT = 60
    Settings.WAIT_HEARTBEATS_CONVERGENCE = 0.0
    Settings.LOG_LEVEL = "WARNING"
    logger.set_level("WARNING")


def wait_convergence(nodes: List, n_neis: int, wait: float = 5, only_direct: bool = False, debug: bool = False) -> None:
    """Block until every node sees ``n_neis`` neighbours, ``AssertionError`` after ``wait`` seconds."""
    deadline = time.monotonic() + wait
    while True:
        counts = [len(n.get_neighbors(only_direct=only_direct)) for n in nodes]
        if all(c == n_neis for c in counts):
            return
        if time.monotonic() > deadline:
            raise AssertionError(f"Convergence not reached: neighbour counts {counts}, wanted {n_neis}")
        time.sleep(0.02)


def full_connection(node, nodes: List) -> None:
    for n in nodes:
        node.connect(n.addr)


def wait_to_finish(nodes: List, timeout: float = 3600, debug: bool = False, poll: float = 0.01) -> None:
    """Block until every node's learning workflow has finished (reference polls every 1 s, utils.py:100-116)."""
    deadline = time.monotonic() + timeout
    while True:
        if all(n.learning_workflow.finished for n in nodes):
            return
        if time.monotonic() > deadline:
            raise TimeoutError("Timeout waiting for nodes to finish")
        time.sleep(poll)


def check_equal_models(nodes: List, atol: float = 1e-1) -> None:
    """Every node holds the same parameters (``np.allclose(atol=1e-1)`` like the reference, utils.py:119-145)."""
    ref = None
    for node in nodes:
        params = node.learner.get_model().get_parameters()
        if ref is None:
            ref = params
            continue
        for a, b in zip(ref, params):
            if not np.allclose(a, b, atol=atol):
                raise AssertionError(f"models differ (max abs diff {float(np.max(np.abs(np.asarray(a, dtype=np.float64) - b)))})")
