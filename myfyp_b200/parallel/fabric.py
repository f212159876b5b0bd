"""The NVLink fabric: device placement, published-slab snapshots, peer access, IPC arenas.

No counterpart in the reference (its "data plane" is pickled bytes through gRPC or a function call,
SURVEY §5.9).  Two deployment shapes:

* **one process, many GPUs** - peer access is enabled between every GPU pair; a :class:`SlabRef`
  carries the tensor itself and kernels dereference it directly (local HBM or over NVLink);
* **one process per GPU** (``torchrun``) - every rank owns an *arena* allocated with ``cudaMalloc``
  through the native library, exports its CUDA-IPC handle once, and maps every peer's arena;
  a :class:`SlabRef` is then ``(rank, offset, generation)`` and resolves to a tensor view of the
  mapped arena.  Published snapshots rotate through ``GENERATIONS`` slots per node so a peer can
  still pull round r while the owner publishes round r+1.

Nodes are *placed* on GPUs round-robin (many per GPU), SURVEY §5.8.
"""

from __future__ import annotations

import dataclasses
import threading
from typing import Optional

import torch

from myfyp_b200.parallel.slab import SlabRef


@dataclasses.dataclass
class Snapshot:
    tensor: Optional[torch.Tensor]
    rank: int = -1
    offset: int = 0
    generation: int = 0


class Fabric:
    _instance: Optional["Fabric"] = None
    _lock = threading.Lock()

    def __init__(self) -> None:
        self.rank = -1  # single-process mode
        self.world = 1
        self._peer_enabled = False
        self._arena = None  # multi-process mode: ArenaPool (see ipc.py)
        self._placement_counter = 0

    @classmethod
    def get(cls) -> "Fabric":
        if cls._instance is None:
            with cls._lock:
                if cls._instance is None:
                    cls._instance = Fabric()
        return cls._instance

    # ------------------------------------------------------------------ placement
    def device_count(self) -> int:
        return torch.cuda.device_count() if torch.cuda.is_available() else 0

    def place(self, index: Optional[int] = None) -> torch.device:
        """Device for the ``index``-th node of this process (round-robin over visible GPUs; CPU without GPUs)."""
        n = self.device_count()
        if n == 0:
            return torch.device("cpu")
        if self.rank >= 0:  # one process per GPU: everything on the local GPU
            return torch.device("cuda", torch.cuda.current_device())
        if index is None:
            index = self._placement_counter
            self._placement_counter += 1
        return torch.device("cuda", index % n)

    # ------------------------------------------------------------------ peer access (single process)
    def enable_peer_access(self) -> int:
        """``cudaDeviceEnablePeerAccess`` among the GPUs this process holds models on (both directions); returns how many ordered
        pairs can reach each other.  GPUs the process does not use are left alone - no context is created on them."""
        return Fabric.ensure_peers(())

    _peers_enabled_for: frozenset = frozenset()

    @classmethod
    def ensure_peers(cls, devices) -> int:
        """Peer access among every GPU in use so far plus ``devices`` (indices).  Cheap when nothing is new."""
        for d in devices:
            cls._gpus_in_use.add(int(d))
        want = frozenset(cls._gpus_in_use)
        if len(want) < 2 or want <= cls._peers_enabled_for:
            return 0
        import ctypes

        from myfyp_b200.ops import native

        lib = native.load()
        arr = (ctypes.c_int * len(want))(*sorted(want))
        ok = int(lib.b2_enable_peer_access(arr, len(want)))
        cls._peers_enabled_for = want
        return ok

    # ------------------------------------------------------------------ snapshots
    def snapshot(self, slab: torch.Tensor) -> Snapshot:
        """Immutable copy of ``slab`` that peers may read until they drop their descriptor."""
        if self._arena is not None and slab.is_cuda:
            return self._arena.publish(slab)
        if slab.is_cuda:
            self.enable_peer_access()
        snap = slab.clone()
        self.make_visible(snap)
        return Snapshot(snap)

    _gpus_in_use: set = set()

    @staticmethod
    def make_visible(t: torch.Tensor) -> None:
        """Call before a descriptor of ``t`` leaves the thread that produced it.  Consumers on the SAME GPU run on that GPU's
        default stream, i.e. behind the kernel that wrote ``t``; a consumer on ANOTHER GPU (peer access) is ordered against
        nothing and would read the allocation's previous contents - typically last round's snapshot.  Once models of this
        process live on more than one GPU the producing stream is therefore drained first (the wait releases the interpreter
        lock); with a single GPU in use nothing is waited for.  (``ops.order_after_producers`` is the consumer-side half.)"""
        if not t.is_cuda:
            return
        Fabric._gpus_in_use.add(t.device.index)
        if len(Fabric._gpus_in_use) > 1:
            torch.cuda.current_stream(t.device).synchronize()

    def has_arena(self) -> bool:
        return self._arena is not None

    def attach_arena(self, arena) -> None:
        self._arena = arena
        self.rank, self.world = arena.rank, arena.world

    def resolve(self, ref: SlabRef) -> torch.Tensor:
        if ref.tensor is not None:
            return ref.tensor
        if self._arena is None:
            raise RuntimeError("descriptor from another process but no IPC arena is attached (Fabric.attach_arena)")
        return self._arena.view(ref.rank, ref.offset, ref.numel, ref.generation)
