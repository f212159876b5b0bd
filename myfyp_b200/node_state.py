"""Per-node mutable state and the rendez-vous primitives stages/commands use.

Reference: p2pfl/node_state.py:49-127.  The attribute names (locks included)
are kept because commands, stages and user tests poke them directly.  In
addition a single condition variable ``cv`` is notified on every state
mutation that somebody may be waiting for, which is what lets the stage
machine be event-driven instead of polling with sleeps.
"""

from __future__ import annotations

import threading
from typing import Dict, List, Optional

from myfyp_b200.experiment import Experiment


class NodeState:
    """State of one federated node."""

    def __init__(self, addr: str, simulation: bool = False) -> None:
        self.addr = addr
        self.status = "Idle"
        self.simulation = simulation
        self.experiment_config = None
        self.experiment: Optional[Experiment] = None
        self.rejoining = False  # entered a running experiment and still waits for its first aggregated model

        # learning bookkeeping
        self.models_aggregated: Dict[str, List[str]] = {}
        self.models_aggregated_next: Dict[str, List[str]] = {}  # announcements of peers that are one round ahead
        self.nei_status: Dict[str, int] = {}
        self.votes_needed: Optional[int] = None   # set by the vote stage while it waits: ballots that complete the box
        if not hasattr(self, "is_direct"):
            self.is_direct = None                 # callable(addr) -> bool, installed by the Node (direct neighbour of ours?); survives clear()
        self.train_set: List[str] = []
        self.train_set_votes: Dict[str, Dict[str, int]] = {}
        self.train_set_votes_next: Dict[str, Dict[str, int]] = {}  # ballots of peers that are already one round ahead
        self.early_models: List[tuple] = []  # model offers that arrived before this node started collecting for that round
        self.early_full_model: Optional[tuple] = None  # (round, source, weights): an aggregated model pushed before we waited for it
        self.aggregate_round = -1  # newest round whose aggregated model this node holds (its own aggregation or an adopted add_model)
        if not hasattr(self, "adopt_early_full_model"):
            self.adopt_early_full_model = None  # set by Node (FullModelCommand.adopt_early)
        if not hasattr(self, "absorb_early_models"):
            self.absorb_early_models = None  # set by Node: files the stashed offers (PartialModelCommand.absorb_early)

        # locks / events (same names as the reference)
        self.train_set_votes_lock = threading.Lock()
        self.start_thread_lock = threading.Lock()
        self.wait_votes_ready_lock = threading.Lock()
        self.model_initialized_lock = threading.Lock()
        self.model_initialized_lock.acquire()
        self.aggregated_model_event = threading.Event()
        self.aggregated_model_event.set()

        # one condition variable for "something changed" wake-ups
        if not hasattr(self, "cv"):
            self.cv = threading.Condition()
            self._version = 0
        self.notify()

    # ------------------------------------------------------------------ wake-ups
    def notify(self) -> None:
        """Wake every thread blocked in :meth:`wait_for`."""
        with self.cv:
            self._version += 1
            self.cv.notify_all()

    def wait_change(self, timeout: Optional[float]) -> bool:
        """Block until the next :meth:`notify` (any state change) or ``timeout``; True if woken by a change."""
        with self.cv:
            v = self._version
            return self.cv.wait_for(lambda: self._version != v, timeout=timeout)

    def version(self) -> int:
        """Change counter: read it BEFORE looking at the state, then :meth:`wait_since` cannot miss a change made in between."""
        return self._version

    def wait_since(self, version: int, timeout: Optional[float]) -> bool:
        """Block until the state has changed since ``version`` was read (or ``timeout``)."""
        with self.cv:
            return self.cv.wait_for(lambda: self._version != version, timeout=timeout)

    def wait_for(self, predicate, timeout: Optional[float]) -> bool:
        """Block until ``predicate()`` is true or ``timeout`` seconds pass."""
        with self.cv:
            return self.cv.wait_for(predicate, timeout=timeout)

    # ------------------------------------------------------------------ experiment
    @property
    def round(self) -> Optional[int]:
        return self.experiment.round if self.experiment is not None else None

    @property
    def total_rounds(self) -> Optional[int]:
        return self.experiment.total_rounds if self.experiment is not None else None

    @property
    def exp_name(self) -> Optional[str]:
        return self.experiment.exp_name if self.experiment is not None else None

    def set_experiment(self, exp_name: str, total_rounds: int, round: int = 0) -> None:
        self.status = "Learning"
        self.experiment = Experiment(exp_name, total_rounds, round)
        self.aggregate_round = round - 1
        self.notify()

    def increase_round(self) -> None:
        if self.experiment is None:
            raise ValueError("Experiment not initialized")
        self.experiment.increase_round()
        self.models_aggregated, self.models_aggregated_next = self.models_aggregated_next, {}
        with self.train_set_votes_lock:
            self.train_set_votes, self.train_set_votes_next = self.train_set_votes_next, {}
        self.notify()

    def clear(self) -> None:
        """Back to the idle state (round becomes None => every stage early-stops)."""
        type(self).__init__(self, self.addr, self.simulation)

    def __str__(self) -> str:
        return (
            f"NodeState(addr={self.addr}, status={self.status}, exp_name={self.exp_name}, round={self.round}, "
            f"total_rounds={self.total_rounds}, simulation={self.simulation}, models_aggregated={self.models_aggregated}, "
            f"nei_status={self.nei_status}, train_set={self.train_set}, train_set_votes={self.train_set_votes})"
        )
