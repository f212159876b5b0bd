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
import os
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
        # one interprocess event per slot: recorded after the publishing copy, waited for by the consumer's stream
        self._lib.b2_ipc_event_create.restype = ctypes.c_void_p
        self._lib.b2_ipc_event_open.restype = ctypes.c_void_p
        with torch.cuda.device(device):
            self._events = [self._lib.b2_ipc_event_create() for _ in range(slots)]
        if not all(self._events):
            native.check(1, "b2_ipc_event_create")
        self._peer_event_handles: Dict[int, List[bytes]] = {}
        self._peer_events: Dict[tuple, int] = {}
        # generation of every slot, in host shared memory every rank of the job maps: a descriptor is only as good as the
        # generation it was issued for (slots rotate)
        self._job = os.environ.get("MASTER_PORT", "0") + "-" + os.environ.get("TORCHELASTIC_RUN_ID", "0")
        self._gen_table = _GenTable(self._job, rank, slots, create=True)
        self._peer_gen: Dict[int, "_GenTable"] = {rank: self._gen_table}
        self._peer_slot_elems: Dict[int, int] = {rank: self.slot_elems}
        self.stale_refused = 0

    def export_handle(self) -> bytes:
        buf = ctypes.create_string_buffer(self.HANDLE_BYTES)
        from myfyp_b200.ops import native

        native.check(self._lib.b2_ipc_export(ctypes.c_void_p(self._base), buf), "b2_ipc_export")
        return bytes(buf.raw)

    def export_event_handles(self) -> List[bytes]:
        from myfyp_b200.ops import native

        out = []
        for ev in self._events:
            buf = ctypes.create_string_buffer(self.HANDLE_BYTES)
            native.check(self._lib.b2_ipc_event_export(ctypes.c_void_p(ev), buf), "b2_ipc_event_export")
            out.append(bytes(buf.raw))
        return out

    def import_handles(self, handles: List[bytes], slot_elems: List[int], slots: List[int], event_handles: Optional[List[List[bytes]]] = None) -> None:
        from myfyp_b200.ops import native

        for r, h in enumerate(handles):
            self._peer_slot_elems[r] = slot_elems[r]
            if event_handles is not None and r != self.rank:
                self._peer_event_handles[r] = event_handles[r]
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

        mine = (self.export_handle(), self.slot_elems, self.slots, self.export_event_handles())
        gathered: List[Any] = [None] * self.world
        dist.all_gather_object(gathered, mine)
        self.import_handles([g[0] for g in gathered], [g[1] for g in gathered], [g[2] for g in gathered], [g[3] for g in gathered])

    def publish(self, slab: torch.Tensor):
        """Copy ``slab`` into the next slot of the ring and return its coordinates.  No host synchronisation: the slot's
        interprocess event is recorded behind the copy, and whoever dereferences the descriptor makes its stream wait for
        that event first (:meth:`view`).  The slot's generation is bumped in the shared table BEFORE the copy is queued, so a
        descriptor of the slot's previous content is refused from this moment on."""
        from myfyp_b200 import ops
        from myfyp_b200.ops import native
        from myfyp_b200.parallel.fabric import Snapshot

        n = slab.numel()
        if n > self.slot_elems:
            raise ValueError(f"slab of {n} elements does not fit an arena slot of {self.slot_elems}")
        with self._lock:
            slot = self._next
            self._next = (self._next + 1) % self.slots
            self._gen += 1
            gen = self._gen
            self._gen_table.set(slot, gen)
        off = slot * self.slot_elems
        view = self._local[off : off + n]
        ops.pull_copy(view, slab)
        with torch.cuda.device(self.device):
            native.check(self._lib.b2_event_record(ctypes.c_void_p(self._events[slot]), native.stream_ptr(self.device)), "b2_event_record")
        return Snapshot(view, self.rank, off, gen)

    def view(self, rank: int, offset: int, numel: int, generation: int = 0) -> torch.Tensor:
        """Tensor view of ``(rank, offset)``.  With a ``generation`` the descriptor is validated against the owner's table
        (:class:`StaleDescriptorError` if the slot has been reused) and the current stream of this GPU is made to wait for the
        slot's publication event, so kernels queued afterwards see the published bytes."""
        from myfyp_b200.learning.exceptions import StaleDescriptorError
        from myfyp_b200.ops import native

        if generation:
            slot = offset // self._peer_slot_elems.get(rank, self.slot_elems)
            tab = self._peer_gen.get(rank)
            if tab is None:
                tab = self._peer_gen[rank] = _GenTable(self._job, rank, 0, create=False)
            now = tab.get(slot)
            if now != generation:
                self.stale_refused += 1
                raise StaleDescriptorError(f"slot {slot} of rank {rank} holds generation {now}, the descriptor was issued for {generation}")
            if rank != self.rank:
                key = (rank, slot)
                ev = self._peer_events.get(key)
                if ev is None and rank in self._peer_event_handles:
                    with torch.cuda.device(self.device):
                        ev = self._lib.b2_ipc_event_open(ctypes.c_char_p(self._peer_event_handles[rank][slot]))
                    if not ev:
                        native.check(1, f"b2_ipc_event_open(rank {rank}, slot {slot})")
                    self._peer_events[key] = ev
                if ev:
                    with torch.cuda.device(self.device):
                        native.check(self._lib.b2_stream_wait_event(native.stream_ptr(self.device), ctypes.c_void_p(ev)), "b2_stream_wait_event")
        return self._peers[rank][offset : offset + numel]

    def still_valid(self, rank: int, offset: int, generation: int) -> bool:
        """Host-side re-check for a consumer that kept a descriptor (or a view) around: has the slot been reused since?"""
        tab = self._peer_gen.get(rank)
        return tab is None or tab.get(offset // self._peer_slot_elems.get(rank, self.slot_elems)) == generation

    def close(self) -> None:
        for r, p in self._peer_bases.items():
            self._lib.b2_ipc_close(ctypes.c_void_p(p))
        self._peer_bases.clear()
        for tab in self._peer_gen.values():
            tab.close()


class _GenTable:
    """``int64[slots]`` in ``/dev/shm`` (one file per rank and job): the generation currently held by every arena slot.
    Written by the owner, mapped read-only by the peers; an aligned 8-byte store / load is atomic on every platform we run on."""

    def __init__(self, job: str, rank: int, slots: int, create: bool) -> None:
        import mmap

        self.path = f"/dev/shm/b2-arena-{job}-{rank}.gen"
        self.owner = create
        if create:
            with open(self.path, "wb") as f:
                f.write(b"\0" * (8 * max(slots, 1)))
        self._f = open(self.path, "r+b" if create else "rb")
        self._mm = mmap.mmap(self._f.fileno(), 0, access=mmap.ACCESS_WRITE if create else mmap.ACCESS_READ)
        self._view = memoryview(self._mm).cast("q")

    def set(self, slot: int, gen: int) -> None:
        self._view[slot] = gen

    def get(self, slot: int) -> int:
        return self._view[slot]

    def close(self) -> None:
        try:
            self._view.release()
            self._mm.close()
            self._f.close()
            if self.owner:
                os.unlink(self.path)
        except Exception:  # noqa: BLE001
            pass


class _Holder:
    """Minimal ``__cuda_array_interface__`` provider so torch can view memory we own."""

    def __init__(self, ptr: int, numel: int) -> None:
        self.__cuda_array_interface__ = {"shape": (numel,), "typestr": "<f4", "data": (ptr, False), "version": 2, "strides": None}


def _wrap(ptr: int, numel: int, device: torch.device) -> torch.Tensor:
    with torch.cuda.device(device):
        return torch.as_tensor(_Holder(ptr, numel), device=device)
