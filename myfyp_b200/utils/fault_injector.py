"""Fault and attack injection for robustness experiments.

The reference only has tests that call ``node.stop()`` mid-run (test/node_test.py:167-199, disabled) and the
fork author's script that poisons one node's ``state_dict`` before learning - sign flip
(exp_SAVE3.txt:60-113) and additive Gaussian noise (:187-234).  Here both are first-class:

* Byzantine learners: ``make_byzantine(node, attack=...)`` wraps ``learner.fit`` so the model the node
  *publishes* each round is poisoned on the device (sign-flip / Gaussian noise / scaled update);
* crash faults: ``FaultInjector.drop_at(node, round, stage)`` stops a node when its workflow reaches a
  stage of a round; ``rejoin_after`` starts a fresh node object at the same address later, which the running
  network pulls back into the experiment (``Settings.ALLOW_REJOIN``).
"""

from __future__ import annotations

import threading
import time
from typing import Callable, List, Optional

import torch


def poison_(slab: torch.Tensor, attack: str, n_trainable: int, reference: Optional[torch.Tensor] = None, scale: float = 1.0, std: float = 1.0, generator: Optional[torch.Generator] = None) -> None:
    """In-place poisoning of the trainable prefix of a flat slab."""
    p = slab[:n_trainable]
    if attack == "sign_flip":
        p.mul_(-abs(scale))
    elif attack == "gaussian":
        p.add_(torch.randn(p.shape, device=p.device, generator=generator) * std)
    elif attack == "scaled_update":
        if reference is None:
            raise ValueError("scaled_update needs the pre-fit parameters")
        p.copy_(reference[:n_trainable] + scale * (p - reference[:n_trainable]))
    else:
        raise ValueError(f"unknown attack {attack!r}")


def make_byzantine(node, attack: str = "sign_flip", scale: float = 1.0, std: float = 1.0) -> None:
    """After every ``fit`` the node's parameters are poisoned before they are published."""
    learner = node.learner
    honest_fit = learner.fit

    def fit():
        model = learner.get_model()
        before = model.slab.clone() if attack == "scaled_update" else None
        out = honest_fit()
        poison_(model.slab, attack, model.layout.n_trainable, before, scale, std)
        return out

    learner.fit = fit  # type: ignore[method-assign]
    node.byzantine = attack


class FaultInjector:
    """Crash / rejoin scheduler driven by the nodes' own workflow histories."""

    def __init__(self) -> None:
        self._threads: List[threading.Thread] = []
        self.events: List[tuple] = []

    def drop_at(self, node, round: int, stage: Optional[str] = "TrainStage", poll: float = 0.005) -> None:
        """Stop ``node`` as soon as it is in ``stage`` of ``round`` (``stage=None``: anywhere in that round)."""
        def watch():
            while True:
                if node.learning_workflow.finished and node.state.round is None and len(node.learning_workflow.history) > 1:
                    return
                if node.state.round == round and node.learning_workflow.history and (stage is None or node.learning_workflow.history[-1] == stage):
                    self.events.append(("drop", node.addr, round, stage, time.time()))
                    node.stop()
                    return
                time.sleep(poll)

        t = threading.Thread(target=watch, daemon=True, name=f"fault-drop-{node.addr}")
        t.start()
        self._threads.append(t)

    def drop_after(self, node, seconds: float) -> None:
        def later():
            time.sleep(seconds)
            self.events.append(("drop", node.addr, None, None, time.time()))
            node.stop()

        t = threading.Thread(target=later, daemon=True, name=f"fault-drop-{node.addr}")
        t.start()
        self._threads.append(t)

    def rejoin_after(self, make_node: Callable[[], object], connect_to: List[str], seconds: float, on_joined: Optional[Callable[[object], None]] = None) -> None:
        def later():
            time.sleep(seconds)
            node = make_node()
            node.start()
            for a in connect_to:
                node.connect(a)
            self.events.append(("rejoin", node.addr, None, None, time.time()))
            if on_joined is not None:
                on_joined(node)

        t = threading.Thread(target=later, daemon=True, name="fault-rejoin")
        t.start()
        self._threads.append(t)

    def rejoin_at(self, make_node: Callable[[], object], connect_to: List[str], watch_node, round: int, poll: float = 0.005, on_joined: Optional[Callable[[object], None]] = None) -> None:
        """Bring a (new) node up and connect it once ``watch_node`` has reached ``round``."""

        def watch():
            while True:
                r = watch_node.state.round
                if r is not None and r >= round:
                    break
                if watch_node.learning_workflow.finished and len(watch_node.learning_workflow.history) > 1:
                    return
                time.sleep(poll)
            node = make_node()
            node.start()
            for a in connect_to:
                node.connect(a)
            self.events.append(("rejoin", node.addr, round, None, time.time()))
            if on_joined is not None:
                on_joined(node)

        t = threading.Thread(target=watch, daemon=True, name="fault-rejoin")
        t.start()
        self._threads.append(t)

    def join(self, timeout: float = 60) -> None:
        for t in self._threads:
            t.join(timeout)
