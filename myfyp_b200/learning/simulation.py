"""Single-box simulation scaling: how the learners of many simulated nodes share the GPUs.

The reference wraps every learner in a ``VirtualNodeLearner`` that ships ``(addr, learner)`` to a Ray actor of a
``SuperActorPool`` sized from the CPUs / GPUs of the Ray cluster (p2pfl/learning/frameworks/simulation/actor_pool.py:39-357,
virtual_learner.py:101-137, utils.py:52-96): learners of different nodes train in parallel *processes*.

Here the unit of parallelism is not a process but a grid dimension:

* nodes are *placed* on GPUs by the fabric (round-robin, many per GPU);
* the learners of one GPU that train the same architecture meet in that GPU's :class:`~myfyp_b200.models.grouped_cnn.CohortTrainer`:
  requests arriving within a few milliseconds are bound to the slots of ONE native plan and run as one CUDA-graph replay per epoch
  (every kernel takes the node index from ``blockIdx.y``) - the counterpart of "submit the job to the pool and block for the result"
  (virtual_learner.py:101-110), with the GPU instead of a worker process doing the batching;
* what remains of the pool itself is admission control (:class:`FitPool`): bound how many learners of one device are inside
  ``fit`` / ``evaluate`` at once (activation memory, fairness between 64 simulated nodes), sized from the device's resources like
  ``pool_size_from_resources``; models outside the node-batched plans (MLP, ResNet-18) run one CUDA graph per node on private streams.

``Settings.DISABLE_RAY`` (the reference's switch) turns the cohort trainer off: every node then trains on its own graph.
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


def pool_size_from_resources(device=None, bytes_per_learner: int = 64 << 20) -> int:
    """How many learners may train on ``device`` at once: free HBM / the activation footprint of one learner, capped at 64 (the
    reference divides the Ray cluster's CPUs / GPUs by the per-client request, simulation/utils.py:52-96)."""
    import torch

    if device is None or not torch.cuda.is_available() or torch.device(device).type != "cuda":
        import os

        return max(1, (os.cpu_count() or 2) // 2)
    free, _ = torch.cuda.mem_get_info(torch.device(device))
    return int(max(1, min(64, free // max(bytes_per_learner, 1))))


class SuperActorPool:
    """Facade with the reference's name (simulation/actor_pool.py:39) over what plays its role here: per-device admission control
    (:class:`FitPool`) plus the per-GPU cohort trainers.  ``stats()`` reports how many fit / evaluate requests were served in how
    many batched launches."""

    @staticmethod
    def configure(slots_per_device: Optional[int] = None, device=None) -> int:
        slots = slots_per_device or pool_size_from_resources(device)
        FitPool.configure(slots)
        return slots

    @staticmethod
    def stats() -> Dict[str, Dict[str, int]]:
        from myfyp_b200.models.grouped_cnn import CohortTrainer

        return {str(k[0]): {"requests": t.batched, "launches": t.batches, "members": len(t._members)} for k, t in CohortTrainer._instances.items()}


def try_init_learner_with_simulation(learner: Learner, slots_per_device: Optional[int] = None) -> Learner:
    """Reference name: ``try_init_learner_with_ray``.  Wraps when simulation is requested."""
    if slots_per_device is not None:
        FitPool.configure(slots_per_device)
    return VirtualNodeLearner(learner)


try_init_learner_with_ray = try_init_learner_with_simulation  # the reference's name
