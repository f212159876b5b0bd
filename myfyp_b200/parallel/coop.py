"""Cooperative FedAvg: when a round's train set is the whole federation, the nodes of a GPU aggregate ONCE, together.

The reference's exchange is pairwise (p2pfl/stages/base_node/train_stage.py:120-176: every member gossips partial aggregates to
every other member, ``gossiper.py:163-239``) and every node then averages what it collected on its own
(``learning/aggregators/fedavg.py:61-69``).  With all members on the NVLink fabric that is ``n (n-1)`` model transfers and ``n``
identical reductions per round for one result.  Here, when every node of the federation is in the train set:

1. the resident nodes of a rank meet in the :class:`CoopHub` after ``fit`` (no model leaves the GPU);
2. one K-way kernel forms the rank's partial ``sum_k n_k theta_k`` (and ``sum_k n_k`` in a spare element) in the collective's buffer;
3. ONE fused kernel per GPU sums the partials of all ranks - in the NVSwitch (``multimem.ld_reduce`` / ``multimem.st``,
   csrc/nvls.cu) - so every rank holds ``sum n_k theta_k`` and ``sum n_k``; with a single rank this step is skipped;
4. every resident node adopts ``sum n_k theta_k / sum n_k``.

Anything irregular - a train set smaller than the federation, a node that does not show up within ``COOP_WINDOW`` seconds, a rank
that never joins the collective (bounded wait inside the kernel), no multicast on the box - and the round takes the reference's
gossip path instead (:meth:`CoopHub.aggregate` returns ``None``); the two paths give the same model (tests/test_node.py).
"""

from __future__ import annotations

import threading
import time
from typing import Dict, List, Optional, Sequence, Tuple

import torch

from myfyp_b200.management.logger import logger


class CoopHub:
    _instance: Optional["CoopHub"] = None

    def __init__(self, rank: int, world: int, device: torch.device, numel: int, resident: Sequence[str], all_nodes: Sequence[str], window: float = 10.0) -> None:
        self.rank, self.world, self.device, self.numel = rank, world, device, numel
        self.resident, self.all_nodes = list(resident), list(all_nodes)
        self.window = window
        self._cv = threading.Condition()
        self._rounds: Dict[Tuple[int, int], dict] = {}
        self._serial: Dict[str, int] = {}   # per node: how many experiments it has started (epochs must grow across experiments)
        self.comm = None
        self.successes = 0
        self.fallbacks = 0
        pad = numel + 64
        if world > 1:
            from myfyp_b200.parallel.collective import NvlsFedAvg, nvls_available

            if device.type == "cuda" and nvls_available():
                self.comm = NvlsFedAvg(pad, rank, world, device)
                self.src, self.out = self.comm.src, self.comm.out
            else:
                self.src = self.out = None  # no multicast: every round falls back
        else:
            self.src = torch.zeros(pad, dtype=torch.float32, device=device)
            self.out = self.src
        self.result = torch.zeros(numel, dtype=torch.float32, device=device) if self.src is not None else None

    # ------------------------------------------------------------------ set-up
    @classmethod
    def configure(cls, rank: int, world: int, device, numel: int, resident: Sequence[str], all_nodes: Sequence[str], window: float = 10.0) -> "CoopHub":
        cls._instance = CoopHub(rank, world, torch.device(device), numel, resident, all_nodes, window)
        return cls._instance

    @classmethod
    def get(cls) -> Optional["CoopHub"]:
        return cls._instance

    @classmethod
    def reset(cls) -> None:
        if cls._instance is not None and cls._instance.comm is not None:
            cls._instance.comm.close()
        cls._instance = None

    def eligible(self, addr: str, train_set: Sequence[str], aggregator) -> bool:
        from myfyp_b200.learning.aggregators.fedavg import FedAvg

        return (
            self.src is not None and type(aggregator) is FedAvg and addr in self.resident and set(train_set) == set(self.all_nodes)
        )

    def experiment_started(self, addr: str) -> None:
        self._serial[addr] = self._serial.get(addr, 0) + 1

    # ------------------------------------------------------------------ one round
    def aggregate(self, addr: str, rnd: int, slab: torch.Tensor, num_samples: int) -> Optional[Tuple[torch.Tensor, int]]:
        """Called by every resident train-set node after its ``fit``.  Returns ``(average, total samples)`` - the same tensor for all
        of them - or ``None`` when the cooperative path did not complete (the caller then gossips as the reference does)."""
        key = (self._serial.get(addr, 0), rnd)
        deadline = time.monotonic() + self.window
        with self._cv:
            st = self._rounds.setdefault(key, {"slabs": {}, "done": False, "result": None, "leader": False})
            st["slabs"][addr] = (slab, int(num_samples))
            self._cv.notify_all()
            while not st["done"]:
                if len(st["slabs"]) >= len(self.resident) and not st["leader"]:
                    st["leader"] = True
                    self._cv.release()
                    try:
                        res = self._run(key, st["slabs"])
                    except Exception as e:  # noqa: BLE001 - the round must go on over the gossip path
                        logger.warning(addr, f"cooperative aggregation failed: {e}")
                        res = None
                    finally:
                        self._cv.acquire()
                    st["result"], st["done"] = res, True
                    self._cv.notify_all()
                    break
                left = deadline - time.monotonic()
                if left <= 0:
                    if not st["leader"]:  # somebody never came: nobody of this rank takes the cooperative path this round
                        st["done"] = True
                        self._cv.notify_all()
                    else:
                        self._cv.wait(0.05)
                    continue
                self._cv.wait(left)
            res = st["result"]
            for k in [k for k in self._rounds if k < key and self._rounds[k]["done"]]:
                del self._rounds[k]
        if res is None:
            self.fallbacks += 1
        return res

    def _run(self, key: Tuple[int, int], slabs: Dict[str, Tuple[torch.Tensor, int]]) -> Optional[Tuple[torch.Tensor, int]]:
        from myfyp_b200 import ops

        n = self.numel
        tensors = [s for s, _ in slabs.values()]
        counts = [float(c) for _, c in slabs.values()]
        cuda = self.device.type == "cuda"
        ctx = torch.cuda.device(self.device) if cuda else _Null()
        with ctx:
            ops.fedavg(tensors, counts, device=self.device, out=self.src[:n])      # sum_k n_k theta_k of this rank, one pass
            self.src[n : n + 4] = torch.tensor([sum(counts), 0.0, 0.0, 0.0], device=self.device)
            if self.world > 1:
                epoch = key[0] * 65536 + key[1] + 1
                self.comm(1.0, epoch=epoch)                                          # in-switch sum over the ranks
            torch.div(self.out[:n], self.out[n], out=self.result)
            total = self.out[n : n + 1].to("cpu", non_blocking=False)                # one small read-back per round and rank
            if cuda:
                torch.cuda.current_stream(self.device).synchronize()
            if self.world > 1 and self.comm.timed_out():
                return None
        self.successes += 1
        return self.result, int(round(float(total[0])))


class _Null:
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False
