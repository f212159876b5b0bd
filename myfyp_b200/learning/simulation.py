"""Single-box simulation scaling: how many learners may train at once.

The reference wraps every learner in a ``VirtualNodeLearner`` that ships ``(addr, learner)`` to a Ray actor of a
``SuperActorPool`` sized from the CPUs / GPUs of the Ray cluster (learning/frameworks/simulation/*.py, SURVEY 2.3,
5.8).  On one B200 box there is nothing to ship: nodes are *placed* on GPUs by the fabric and their epochs are CUDA
graphs on private streams, so many nodes per GPU already run concurrently.  What remains useful from the pool is the
admission control - bound how many learners of one GPU are inside ``fit`` / ``evaluate`` at the same time (memory for
activations, fairness between 64 simulated nodes) - and the same entry point name so scripts keep working.
"""

from __future__ import annotations

import threading
from typing import Dict, Optional

from myfyp_b200.learning.learner import Learner


class FitPool:
    """Per-device semaphores: at most ``slots`` learners of one device train at once (the SuperActorPool analogue)."""

    _lock = threading.Lock()
    _sems: Dict[str, threading.BoundedSemaphore] = {}
    slots_per_device: int = 16

    @classmethod
    def semaphore(cls, device) -> threading.BoundedSemaphore:
        key = str(device)
        with cls._lock:
            if key not in cls._sems:
                cls._sems[key] = threading.BoundedSemaphore(cls.slots_per_device)
            return cls._sems[key]

    @classmethod
    def configure(cls, slots_per_device: int) -> None:
        with cls._lock:
            cls.slots_per_device = slots_per_device
            cls._sems.clear()


class VirtualNodeLearner(Learner):
    """Delegates to a real learner, queueing ``fit`` / ``evaluate`` behind the device's :class:`FitPool`."""

    def __init__(self, learner: Learner, addr: Optional[str] = None) -> None:  # noqa: D401 - wraps, does not call super().__init__
        self.learner = learner
        self.addr = addr or learner._self_addr

    def __getattr__(self, name):  # everything not overridden goes to the wrapped learner
        return getattr(self.__dict__["learner"], name)

    def _device(self):
        return getattr(self.learner.get_model(), "device", "cpu")

    def set_addr(self, addr: str) -> None:
        self.addr = addr
        self.learner.set_addr(addr)

    def set_model(self, model) -> None:
        self.learner.set_model(model)

    def get_model(self):
        return self.learner.get_model()

    def set_data(self, data) -> None:
        self.learner.set_data(data)

    def get_data(self):
        return self.learner.get_data()

    def set_epochs(self, epochs: int) -> None:
        self.learner.set_epochs(epochs)

    def fit(self):
        with FitPool.semaphore(self._device()):
            return self.learner.fit()

    def evaluate(self):
        with FitPool.semaphore(self._device()):
            return self.learner.evaluate()

    def interrupt_fit(self) -> None:  # the reference's VirtualNodeLearner raises NotImplementedError here
        self.learner.interrupt_fit()

    def get_framework(self) -> str:
        return self.learner.get_framework()


def try_init_learner_with_simulation(learner: Learner, slots_per_device: Optional[int] = None) -> Learner:
    """Reference name: ``try_init_learner_with_ray``.  Wraps when simulation is requested."""
    if slots_per_device is not None:
        FitPool.configure(slots_per_device)
    return VirtualNodeLearner(learner)


try_init_learner_with_ray = try_init_learner_with_simulation  # the reference's name
