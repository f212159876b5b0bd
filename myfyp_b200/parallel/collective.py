"""Fused compute + collective kernels across the GPUs of one box (no NCCL on these paths).

``TwoShotFedAvg``: sample-weighted average of one slab per rank, as ONE kernel per GPU that reduce-scatters by coordinate
shard and all-gathers the reduced shards with peer loads, synchronised by flags in peer-mapped memory (kernel A2,
``csrc/collective.cu``).  Every rank owns a *symmetric region* ``[flag block | input slab | output slab]`` allocated with
``cudaMalloc`` through the native library and exported once over CUDA IPC.

Two ways to run it:

* ``TwoShotFedAvg(n, rank, world, device)`` under ``torchrun`` - one rank per GPU; handles travel through
  ``torch.distributed.all_gather_object`` at construction (plumbing), nothing but the kernel runs per call;
* ``TwoShotFedAvg.emulated(n, world, device)`` - all ranks in one launch on one GPU (every pointer local).  Kernels that wait
  on each other must be co-resident, which separate launches on one GPU do not guarantee; this is how the kernel is tested
  on a single-GPU box.

The NCCL path (``nccl_fedavg``) is the *baseline* the sweep benchmark compares against.
"""

from __future__ import annotations

import ctypes
from typing import List, Optional, Sequence

import torch

from myfyp_b200.ops import native
from myfyp_b200.parallel.ipc import _wrap

MAX_RANKS = 16


class _Args(ctypes.Structure):
    _fields_ = [
        ("srcs", ctypes.c_void_p * MAX_RANKS), ("outs", ctypes.c_void_p * MAX_RANKS), ("flags", ctypes.c_void_p * MAX_RANKS),
        ("n", ctypes.c_int64), ("world", ctypes.c_int), ("rank", ctypes.c_int),
    ]


def _region_bytes(n: int, flag_bytes: int) -> int:
    return _pad(flag_bytes) + 2 * 4 * n


def _pad(b: int) -> int:
    return (b + 255) // 256 * 256


class TwoShotFedAvg:
    def __init__(self, n: int, rank: int, world: int, device: torch.device, _emulated: bool = False) -> None:
        if n % 4:
            raise ValueError("n must be a multiple of 4 (slabs are padded to 64 elements)")
        if world > MAX_RANKS:
            raise ValueError(f"at most {MAX_RANKS} ranks")
        self.n, self.rank, self.world, self.device = n, rank, world, device
        self.lib = native.load()
        self.lib.b2_arena_alloc.restype = ctypes.c_void_p
        self.lib.b2_ipc_open.restype = ctypes.c_void_p
        self.lib.b2_twoshot_flag_bytes.restype = ctypes.c_int64
        self.flag_bytes = int(self.lib.b2_twoshot_flag_bytes())
        self.epoch = 0
        self._emulated = _emulated
        self._bases: List[int] = []
        self._opened: List[int] = []
        nbytes = _region_bytes(n, self.flag_bytes)
        with torch.cuda.device(device):
            if _emulated:
                self._bases = [int(self.lib.b2_arena_alloc(ctypes.c_int64(nbytes)) or 0) for _ in range(world)]
            else:
                import torch.distributed as dist

                mine = int(self.lib.b2_arena_alloc(ctypes.c_int64(nbytes)) or 0)
                if not mine:
                    native.check(1, "b2_arena_alloc")
                buf = ctypes.create_string_buffer(64)
                native.check(self.lib.b2_ipc_export(ctypes.c_void_p(mine), buf), "b2_ipc_export")
                handles: List[Optional[bytes]] = [None] * world
                dist.all_gather_object(handles, bytes(buf.raw))
                for g, h in enumerate(handles):
                    if g == rank:
                        self._bases.append(mine)
                    else:
                        p = int(self.lib.b2_ipc_open(ctypes.c_char_p(h)) or 0)
                        if not p:
                            native.check(1, f"b2_ipc_open(rank {g})")
                        self._bases.append(p)
                        self._opened.append(p)
        if any(b == 0 for b in self._bases):
            native.check(1, "b2_arena_alloc")
        off_src, off_out = _pad(self.flag_bytes), _pad(self.flag_bytes) + 4 * n
        self.srcs = [_wrap(b + off_src, n, device) for b in self._bases]
        self.outs = [_wrap(b + off_out, n, device) for b in self._bases]
        self._off = (off_src, off_out)
        # the address table never changes: upload it once; weights and epoch travel as kernel arguments
        ranks = range(world) if _emulated else [rank]
        blob = b"".join(bytes(self._fill(r)) for r in ranks)
        self._args_dev = torch.frombuffer(bytearray(blob), dtype=torch.uint8).to(device)
        self._nranks = len(list(ranks))
        torch.cuda.synchronize(device)

    @classmethod
    def emulated(cls, n: int, world: int, device: torch.device) -> "TwoShotFedAvg":
        return cls(n, 0, world, device, _emulated=True)

    @property
    def src(self) -> torch.Tensor:
        """This rank's input slab (write the model here, or train directly in it)."""
        return self.srcs[self.rank]

    @property
    def out(self) -> torch.Tensor:
        return self.outs[self.rank]

    def _fill(self, rank: int) -> _Args:
        a = _Args()
        for g in range(self.world):
            a.srcs[g] = self._bases[g] + self._off[0]
            a.outs[g] = self._bases[g] + self._off[1]
            a.flags[g] = self._bases[g]
        a.n, a.world, a.rank = self.n, self.world, rank
        return a

    def __call__(self, weights: Sequence[float], max_ctas: int = 0) -> torch.Tensor:
        """Launch the fused kernel on the current stream; returns this rank's output slab (all ranks' in emulated mode: ``outs``)."""
        self.epoch += 1
        with native.device_guard(self.device):
            native.check(
                self.lib.b2_fedavg_twoshot(
                    native.ptr(self._args_dev), self._nranks, ctypes.c_int64(self.n), native.float_array(weights), self.world, ctypes.c_uint(self.epoch), max_ctas,
                    native.stream_ptr(self.device),
                ),
                "b2_fedavg_twoshot",
            )
        native.count_launch()
        return self.out

    def close(self) -> None:
        for p in self._opened:
            self.lib.b2_ipc_close(ctypes.c_void_p(p))
        self._opened.clear()


def nccl_fedavg(x: torch.Tensor, n_k: float, total: float) -> torch.Tensor:
    """BASELINE: pre-weight, NCCL all-reduce, done (what a torch.distributed FedAvg would do)."""
    import torch.distributed as dist

    y = x * (n_k / total)
    dist.all_reduce(y)
    return y
