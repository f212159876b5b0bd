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


# ---------------------------------------------------------------------------------------------- slab-shaped side information
def layers_from_flat(flat: Any, layout: SlabLayout) -> List[Any]:
    """A slab-shaped flat vector (Scaffold's delta_y_i / delta_c_i / global_c) as the reference's wire form: one NumPy array
    per ``state_dict`` entry, in ``state_dict`` order, padding dropped (reference: pytorch/callbacks/scaffold_callback.py:131-135).
    A vector that only covers the trainable prefix yields zeros for the buffers behind it."""
    import numpy as np

    host = flat.detach().cpu().numpy() if isinstance(flat, torch.Tensor) else np.asarray(flat)
    host = host.reshape(-1)
    out = []
    for e in layout.in_order():
        if e.offset + e.numel <= host.size:
            out.append(host[e.offset : e.offset + e.numel].reshape(e.shape).astype(np.float32, copy=True))
        else:
            out.append(np.zeros(e.shape, dtype=np.float32))
    return out


def flat_from_wire(value: Any, layout: SlabLayout, device: Any) -> torch.Tensor:
    """The inverse: a flat fp32 tensor of ``layout.total`` elements on ``device`` from whatever form the value travelled in -
    a tensor (fabric form), a 1-D array of slab length (flat form), or the reference's list of per-layer arrays.
    Anything else is a malformed message and raises instead of being broadcast over whole layers."""
    import numpy as np

    if isinstance(value, torch.Tensor):
        t = value.reshape(-1)
        if t.numel() not in (layout.total, layout.n_trainable):
            raise ValueError(f"flat side information has {t.numel()} elements, the slab has {layout.total}")
        if t.numel() < layout.total:
            full = torch.zeros(layout.total, dtype=torch.float32, device=t.device)
            full[: t.numel()] = t
            t = full
        return t if str(t.device) == str(torch.device(device)) else t.to(device)
    if isinstance(value, np.ndarray) and value.ndim == 1 and value.size in (layout.total, layout.n_trainable) and value.dtype != object:
        host = np.zeros(layout.total, dtype=np.float32)
        host[: value.size] = value
        return torch.from_numpy(host).to(device)
    entries = layout.in_order()
    if not isinstance(value, (list, tuple)) or len(value) != len(entries):
        got = f"{len(value)} items" if isinstance(value, (list, tuple)) else f"{type(value).__name__} of shape {getattr(value, 'shape', '?')}"
        raise ValueError(f"side information must be a slab-length vector or a list of {len(entries)} per-layer arrays (got {got})")
    host = np.zeros(layout.total, dtype=np.float32)
    for e, layer in zip(entries, value):
        a = np.asarray(layer, dtype=np.float32).reshape(-1)
        if a.size != e.numel:
            raise ValueError(f"layer {e.name}: expected {e.numel} values, got {a.size}")
        host[e.offset : e.offset + e.numel] = a
    return torch.from_numpy(host).to(device)
