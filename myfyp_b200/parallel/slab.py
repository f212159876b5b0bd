"""Flat parameter slabs and the descriptors that travel in weights messages.

A model's ``state_dict`` lives in ONE flat fp32 buffer (the *slab*): trainable
parameters first, then buffers, every entry starting on a 256-byte boundary so each
view is vector-load aligned.  Aggregation, diffusion, the optimizer and the Scaffold /
FedProx terms are then single kernels over contiguous memory, and "sending a model"
to a peer on the NVLink fabric means sending a :class:`SlabRef`.

The reference has no counterpart: it ships ``[ndarray...]`` pickles
(p2pfl/learning/frameworks/p2pfl_model.py:71-85) rebuilt from ``state_dict()`` on every call.
"""

from __future__ import annotations

import dataclasses
import hashlib
from typing import Any, Dict, List, Optional, Tuple

import torch

ALIGN_ELEMS = 64  # 64 fp32 = 256 B


@dataclasses.dataclass(frozen=True)
class SlabEntry:
    name: str
    shape: Tuple[int, ...]
    dtype: torch.dtype
    offset: int  # in fp32 elements
    numel: int
    trainable: bool


class SlabLayout:
    """Where every ``state_dict`` entry sits in the flat buffer."""

    def __init__(self, entries: List[SlabEntry], order: List[str], n_trainable: int, total: int) -> None:
        self.entries = {e.name: e for e in entries}
        self.order = order  # state_dict order = wire order of the reference's parameter list
        self.n_trainable = n_trainable  # length of the trainable prefix (padded)
        self.total = total
        sig = ";".join(f"{n}:{tuple(self.entries[n].shape)}" for n in order)
        self.hash = int.from_bytes(hashlib.sha1(sig.encode()).digest()[:8], "little") >> 1

    @staticmethod
    def from_module(module: torch.nn.Module) -> "SlabLayout":
        sd = module.state_dict(keep_vars=True)
        param_names = {n for n, _ in module.named_parameters()}
        order = list(sd.keys())
        entries: List[SlabEntry] = []
        off = 0
        for trainable in (True, False):
            for n in order:
                t = sd[n]
                is_tr = n in param_names and t.requires_grad and t.is_floating_point()
                if is_tr != trainable:
                    continue
                entries.append(SlabEntry(n, tuple(t.shape), t.dtype, off, t.numel(), is_tr))
                off += (t.numel() + ALIGN_ELEMS - 1) // ALIGN_ELEMS * ALIGN_ELEMS
            if trainable:
                n_trainable = off
        return SlabLayout(entries, order, n_trainable, max(off, ALIGN_ELEMS))

    def in_order(self) -> List[SlabEntry]:
        return [self.entries[n] for n in self.order]


@dataclasses.dataclass
class SlabRef:
    """Descriptor of a published (immutable for its round) parameter slab.

    In one process ``tensor`` is the slab itself (the reference keeps the memory alive).
    Across processes ``tensor`` is None and ``(rank, offset)`` address the owner's
    IPC-mapped arena (see :mod:`myfyp_b200.parallel.fabric`).
    """

    tensor: Optional[torch.Tensor]
    numel: int
    layout_hash: int
    device: str
    additional_info: Dict[str, Any] = dataclasses.field(default_factory=dict)
    rank: int = -1
    offset: int = 0
    generation: int = 0

    def resolve(self) -> torch.Tensor:
        if self.tensor is not None:
            return self.tensor
        from myfyp_b200.parallel.fabric import Fabric

        return Fabric.get().resolve(self)

    def nbytes(self) -> int:
        return 4 * self.numel
