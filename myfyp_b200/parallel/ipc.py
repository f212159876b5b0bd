"""Cross-process slab sharing: CUDA-IPC arenas and wire form of descriptors.

One process per GPU (``torchrun``).  Each rank allocates ONE arena with ``cudaMalloc`` through the native
library, exports its ``cudaIpcMemHandle_t`` once, and maps all peers' arenas (``cudaIpcOpenMemHandle``), so a
published slab is addressable on every GPU of the box as ``(rank, offset)`` - the same load instructions the
single-process path uses, now on IPC-mapped peer pointers over NVLink.

Arena layout: ``slots`` equal slots of ``slot_elems`` fp32; ``publish`` copies a slab into the next slot
(round-robin).  With ``slots >= 3 x resident nodes`` a snapshot survives at least two further publishes of
every resident node, i.e. longer than any peer may lawfully still be reading it (a peer acts on round-r
descriptors only while it is itself in round r, and no node runs more than one round ahead of a neighbour).
"""

from __future__ import annotations

import ctypes
import pickle
import threading
from typing import Any, Dict, List

import torch

from myfyp_b200.parallel.slab import SlabRef

_MAGIC = b"B2SLABREF\x00"
_addr_to_rank: Dict[str, int] = {}


def is_descriptor(data: bytes) -> bool:
    return data[: len(_MAGIC)] == _MAGIC


def encode_descriptor(ref: Any) -> bytes:
    """Wire form of a :class:`SlabRef` for a peer process on the same box (NumPy-safe additional info)."""
    if not isinstance(ref, SlabRef):
        raise TypeError(f"cannot put {type(ref).__name__} on the wire")
    if ref.rank < 0:
        raise ValueError("descriptor is process-local (no IPC arena attached); send bytes instead")
    from myfyp_b200.learning.p2pfl_model import _info_to_numpy

    info = {k: (v if not isinstance(v, dict) else {kk: _wire_info(vv) for kk, vv in v.items()}) for k, v in ref.additional_info.items()}
    return _MAGIC + pickle.dumps((ref.rank, ref.offset, ref.numel, ref.layout_hash, ref.generation, ref.device, _info_to_numpy(info)))


def _wire_info(v: Any) -> Any:
    """Flat device slabs inside additional_info (Scaffold deltas) travel as nested descriptors when they live in an arena."""
    return v


def decode_descriptor(data: bytes) -> SlabRef:
    rank, offset, numel, layout_hash, generation, device, info = pickle.loads(data[len(_MAGIC) :])
    return SlabRef(None, numel, layout_hash, device, info, rank, offset, generation)


def register_addresses(mapping: Dict[str, int]) -> None:
    """``node address -> owning rank`` for every node of the job (exchanged once at start-up)."""
    _addr_to_rank.update(mapping)


def address_on_fabric(addr: str) -> bool:
    from myfyp_b200.parallel.fabric import Fabric

    return Fabric.get().has_arena() and addr in _addr_to_rank


class ArenaPool:
    """This rank's publish arena + read-only views of every peer's arena."""

    HANDLE_BYTES = 64

    def __init__(self, rank: int, world: int, device: torch.device, slot_elems: int, slots: int) -> None:
        from myfyp_b200.ops import native

        self.rank, self.world, self.device = rank, world, device
        self.slot_elems = (slot_elems + 63) // 64 * 64
        self.slots = slots
        self._next = 0
        self._gen = 0
        self._lock = threading.Lock()
        self._lib = native.load()
        self._lib.b2_arena_alloc.restype = ctypes.c_void_p
        self._lib.b2_ipc_open.restype = ctypes.c_void_p
        nbytes = 4 * self.slot_elems * slots
        with torch.cuda.device(device):
            base = self._lib.b2_arena_alloc(ctypes.c_int64(nbytes))
        if not base:
            native.check(1, "b2_arena_alloc")
        self._base = int(base)
        self._nbytes = nbytes
        self._local = _wrap(self._base, self.slot_elems * slots, device)
        self._peers: Dict[int, torch.Tensor] = {rank: self._local}
        self._peer_bases: Dict[int, int] = {}

    def export_handle(self) -> bytes:
        buf = ctypes.create_string_buffer(self.HANDLE_BYTES)
        from myfyp_b200.ops import native

        native.check(self._lib.b2_ipc_export(ctypes.c_void_p(self._base), buf), "b2_ipc_export")
        return bytes(buf.raw)

    def import_handles(self, handles: List[bytes], slot_elems: List[int], slots: List[int]) -> None:
        from myfyp_b200.ops import native

        for r, h in enumerate(handles):
            if r == self.rank:
                continue
            with torch.cuda.device(self.device):
                p = self._lib.b2_ipc_open(ctypes.c_char_p(h))
            if not p:
                native.check(1, f"b2_ipc_open(rank {r})")
            self._peer_bases[r] = int(p)
            self._peers[r] = _wrap(int(p), slot_elems[r] * slots[r], self.device)

    def exchange(self) -> None:
        """All-gather the IPC handles over ``torch.distributed`` (object collective: plumbing, not the hot path)."""
        import torch.distributed as dist

        mine = (self.export_handle(), self.slot_elems, self.slots)
        gathered: List[Any] = [None] * self.world
        dist.all_gather_object(gathered, mine)
        self.import_handles([g[0] for g in gathered], [g[1] for g in gathered], [g[2] for g in gathered])

    def publish(self, slab: torch.Tensor):
        from myfyp_b200 import ops
        from myfyp_b200.parallel.fabric import Snapshot

        n = slab.numel()
        if n > self.slot_elems:
            raise ValueError(f"slab of {n} elements does not fit an arena slot of {self.slot_elems}")
        with self._lock:
            slot = self._next
            self._next = (self._next + 1) % self.slots
            self._gen += 1
            gen = self._gen
        off = slot * self.slot_elems
        view = self._local[off : off + n]
        ops.pull_copy(view, slab)
        # the descriptor is sent by the host right after: make the copy visible before any peer can dereference it
        torch.cuda.current_stream(self.device).synchronize()
        return Snapshot(view, self.rank, off, gen)

    def view(self, rank: int, offset: int, numel: int) -> torch.Tensor:
        return self._peers[rank][offset : offset + numel]

    def close(self) -> None:
        for r, p in self._peer_bases.items():
            self._lib.b2_ipc_close(ctypes.c_void_p(p))
        self._peer_bases.clear()


class _Holder:
    """Minimal ``__cuda_array_interface__`` provider so torch can view memory we own."""

    def __init__(self, ptr: int, numel: int) -> None:
        self.__cuda_array_interface__ = {"shape": (numel,), "typestr": "<f4", "data": (ptr, False), "version": 2, "strides": None}


def _wrap(ptr: int, numel: int, device: torch.device) -> torch.Tensor:
    with torch.cuda.device(device):
        return torch.as_tensor(_Holder(ptr, numel), device=device)
