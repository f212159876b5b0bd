"""Model wrapper: ``P2PFLModel`` (framework-agnostic contract) and the torch/slab binding.

Contract and wire format follow the reference
(p2pfl/learning/frameworks/p2pfl_model.py:30-195, pytorch/lightning_model.py:37-108):
``encode_parameters()`` emits ``pickle({"params": [fp32 ndarray per state_dict entry],
"additional_info": {...}})`` so a reference node can decode what we send and vice versa.

B200-first differences:

* parameters live in a flat device slab (:mod:`myfyp_b200.parallel.slab`); the wire
  pickle is only materialised for off-fabric peers, checkpoints and tests;
* ``build_copy`` never deep-copies the ``nn.Module`` (reference K7, p2pfl_model.py:185):
  a received or aggregated model is a *held* model = a flat slab + the shared layout;
* ``set_parameters`` also accepts a :class:`SlabRef` (device descriptor) or another model.
"""

from __future__ import annotations

import copy
import pickle
from typing import Any, Dict, List, Optional, Tuple, Union

import numpy as np
import torch

from myfyp_b200.learning.exceptions import DecodingParamsError, ModelNotMatchingError
from myfyp_b200.parallel.slab import SlabLayout, SlabRef


class P2PFLModel:
    """Framework-agnostic model contract (abstract parameter access)."""

    def __init__(
        self,
        model: Any,
        params: Optional[Union[List[np.ndarray], bytes]] = None,
        num_samples: Optional[int] = None,
        contributors: Optional[List[str]] = None,
        additional_info: Optional[Dict[str, Any]] = None,
    ) -> None:
        self.model = model
        self.contributors: List[str] = list(contributors) if contributors is not None else []
        self.num_samples: int = num_samples if num_samples is not None else 0
        self.additional_info: Dict[str, Any] = dict(additional_info) if additional_info is not None else {}
        if params is not None:
            self.set_parameters(params)

    def get_model(self) -> Any:
        return self.model

    # -- wire format ---------------------------------------------------------------
    def encode_parameters(self, params: Optional[List[np.ndarray]] = None) -> bytes:
        if params is None:
            params = self.get_parameters()
        return pickle.dumps({"params": params, "additional_info": _info_to_numpy(self.additional_info)})

    def decode_parameters(self, data: bytes) -> Tuple[List[np.ndarray], Dict[str, Any]]:
        try:
            loaded = pickle.loads(data)
            return loaded["params"], loaded["additional_info"]
        except Exception as e:  # noqa: BLE001
            raise DecodingParamsError("Error decoding parameters") from e

    def get_parameters(self) -> List[np.ndarray]:
        raise NotImplementedError

    def set_parameters(self, params: Union[List[np.ndarray], bytes]) -> None:
        raise NotImplementedError

    # -- callback side-channel -----------------------------------------------------------
    def add_info(self, callback: str, info: Any) -> None:
        self.additional_info[callback] = info

    def get_info(self, callback: Optional[str] = None) -> Any:
        if callback is None:
            return self.additional_info
        return self.additional_info[callback]

    # -- contribution ----------------------------------------------------------------------
    def set_contribution(self, contributors: List[str], num_samples: int) -> None:
        self.contributors = list(contributors)
        self.num_samples = num_samples

    def get_contributors(self) -> List[str]:
        if self.contributors == []:
            raise ValueError("Contributors are empty")
        return self.contributors

    def get_num_samples(self) -> int:
        if self.num_samples == 0:
            raise ValueError("Number of samples required")
        return self.num_samples

    def build_copy(self, **kwargs) -> "P2PFLModel":
        return self.__class__(copy.deepcopy(self.model), **kwargs)

    def snapshot_model(self) -> "P2PFLModel":
        """The object handed to the aggregator after ``fit`` (slab models return an immutable copy)."""
        return self

    def get_framework(self) -> str:
        raise NotImplementedError


def _info_to_numpy(info: Any) -> Any:
    """Device tensors inside additional_info become NumPy so the pickle is portable."""
    if isinstance(info, torch.Tensor):
        return info.detach().cpu().numpy()
    if isinstance(info, dict):
        return {k: _info_to_numpy(v) for k, v in info.items()}
    if isinstance(info, (list, tuple)):
        return type(info)(_info_to_numpy(v) for v in info)
    return info


def _slab_vectors_to_layers(info: Any, layout: SlabLayout) -> Any:
    from myfyp_b200.parallel.slab import layers_from_flat

    if isinstance(info, torch.Tensor) and info.dim() == 1 and info.numel() in (layout.total, layout.n_trainable) and info.numel() > 0:
        return layers_from_flat(info, layout)
    if isinstance(info, dict):
        return {k: _slab_vectors_to_layers(v, layout) for k, v in info.items()}
    return info


class TorchP2PFLModel(P2PFLModel):
    """PyTorch model whose ``state_dict`` is backed by one flat fp32 slab.

    ``model`` may be ``None`` for a *held* model (a received / aggregated parameter set):
    then ``layout`` and ``slab`` must be given and no ``nn.Module`` exists at all.
    """

    def __init__(
        self,
        model: Optional[torch.nn.Module],
        params: Optional[Union[List[np.ndarray], bytes, SlabRef, torch.Tensor]] = None,
        num_samples: Optional[int] = None,
        contributors: Optional[List[str]] = None,
        additional_info: Optional[Dict[str, Any]] = None,
        layout: Optional[SlabLayout] = None,
        device: Optional[Union[str, torch.device]] = None,
        slab: Optional[torch.Tensor] = None,
    ) -> None:
        self.model = model
        self.layout: SlabLayout
        self.slab: torch.Tensor
        self.grad_slab: Optional[torch.Tensor] = None
        self._int_entries: List[str] = []
        self._ref: Optional[SlabRef] = None  # held models: the descriptor (arena coordinates) the slab came with
        if model is not None:
            if device is not None:
                model.to(device)
            self.layout = SlabLayout.from_module(model)
            self._adopt_module(model)
        else:
            if layout is None:
                raise ValueError("a held model needs a layout")
            self.layout = layout
            dev = torch.device(device) if device is not None else (slab.device if slab is not None else torch.device("cpu"))
            self.slab = slab if slab is not None else torch.zeros(layout.total, dtype=torch.float32, device=dev)
        super().__init__(self.model, params, num_samples, contributors, additional_info)

    # ------------------------------------------------------------------ slab plumbing
    @property
    def device(self) -> torch.device:
        return self.slab.device

    def _adopt_module(self, module: torch.nn.Module) -> None:
        """Move every state_dict entry into the slab and re-point the module at the views."""
        sd = module.state_dict(keep_vars=True)
        dev = next(iter(sd.values())).device if len(sd) else torch.device("cpu")
        slab = torch.zeros(self.layout.total, dtype=torch.float32, device=dev)
        grad = torch.zeros(max(self.layout.n_trainable, 1), dtype=torch.float32, device=dev)
        for e in self.layout.in_order():
            t = sd[e.name]
            view = slab[e.offset : e.offset + e.numel].view(e.shape)
            if t.dtype == torch.float32:
                view.copy_(t.detach())
                t.data = view
                if e.trainable:
                    t.grad = grad[e.offset : e.offset + e.numel].view(e.shape)
            else:
                # integer / half buffers keep their own storage; the slab carries an fp32 mirror
                view.copy_(t.detach().to(torch.float32))
                self._int_entries.append(e.name)
        self.slab, self.grad_slab = slab, grad
        if slab.is_cuda:  # the fabric learns which GPUs of this process hold models (see Fabric.make_visible)
            from myfyp_b200.parallel.fabric import Fabric

            Fabric.ensure_peers((slab.device.index,))  # peer access among the GPUs that hold models - and only those

    def sync_mirrors_to_slab(self) -> None:
        """Refresh the fp32 mirrors of non-fp32 buffers (e.g. BatchNorm num_batches_tracked)."""
        if self.model is None or not self._int_entries:
            return
        sd = self.model.state_dict(keep_vars=True)
        for n in self._int_entries:
            e = self.layout.entries[n]
            self.slab[e.offset : e.offset + e.numel].view(e.shape).copy_(sd[n].detach().to(torch.float32))

    def sync_mirrors_from_slab(self) -> None:
        """"Average then cast" semantics for non-fp32 entries."""
        if self.model is None or not self._int_entries:
            return
        sd = self.model.state_dict(keep_vars=True)
        for n in self._int_entries:
            e = self.layout.entries[n]
            v = self.slab[e.offset : e.offset + e.numel].view(e.shape)
            sd[n].data.copy_(v.round().to(sd[n].dtype) if not sd[n].dtype.is_floating_point else v.to(sd[n].dtype))

    def to(self, device: Union[str, torch.device]) -> "TorchP2PFLModel":
        device = torch.device(device)
        if device == self.slab.device:
            return self
        if self.model is not None:
            self.sync_mirrors_to_slab()
            # detach views, move, re-adopt
            sd = self.model.state_dict(keep_vars=True)
            for e in self.layout.in_order():
                if sd[e.name].dtype == torch.float32:
                    sd[e.name].data = sd[e.name].data.clone()
                    sd[e.name].grad = None
            self.model.to(device)
            self._int_entries = []
            self._adopt_module(self.model)
            if device.type == "cuda" and (getattr(self.model, "native", None) is not None or getattr(self.model, "wants_native", False)):
                if hasattr(self.model, "features"):
                    from myfyp_b200.models.native_cnn import attach_native_cnn

                    attach_native_cnn(self, getattr(self.model, "image_size", 32))
                elif hasattr(self.model, "layers") and hasattr(self.model, "bn1"):
                    from myfyp_b200.models.native_resnet import attach_native_resnet

                    attach_native_resnet(self, getattr(self.model, "image_size", 32))
                else:
                    from myfyp_b200.models.native_mlp import attach_native_mlp

                    attach_native_mlp(self)
        else:
            self.slab = self.slab.to(device)
        return self

    # ------------------------------------------------------------------ publication (data plane)
    def publish(self) -> SlabRef:
        """Immutable snapshot of the current parameters as a device descriptor.

        One device-to-device copy (4N bytes).  The owner may keep training while peers
        pull the snapshot (the "double-buffered published slab" of SURVEY §7.1).
        """
        self.sync_mirrors_to_slab()
        from myfyp_b200.parallel.fabric import Fabric

        snap = Fabric.get().snapshot(self.slab)
        return SlabRef(snap.tensor, self.layout.total, self.layout.hash, str(self.slab.device), dict(self.additional_info), snap.rank, snap.offset, snap.generation)

    def as_ref(self) -> SlabRef:
        """Descriptor of the slab itself - only for held models nobody mutates.

        A slab that arrived as a descriptor keeps that descriptor (it may name an IPC arena slot that peers in
        other processes can map).  With an arena attached, a process-local slab (an aggregation result) is copied
        into an arena slot once, because a peer process can only dereference arena coordinates.
        """
        import dataclasses

        if self._ref is not None:
            return dataclasses.replace(self._ref, additional_info=dict(self.additional_info))
        from myfyp_b200.parallel.fabric import Fabric

        if Fabric.get().has_arena() and self.slab.is_cuda:
            snap = Fabric.get().snapshot(self.slab)
            self._ref = SlabRef(snap.tensor, self.layout.total, self.layout.hash, str(self.slab.device), {}, snap.rank, snap.offset, snap.generation)
            return dataclasses.replace(self._ref, additional_info=dict(self.additional_info))
        if not getattr(self, "_visible", False):
            Fabric.make_visible(self.slab)  # e.g. an aggregation result still being written on this GPU's stream
            self._visible = True
        return SlabRef(self.slab, self.layout.total, self.layout.hash, str(self.slab.device), dict(self.additional_info))

    def encode_for_transport(self, device_ok: bool) -> Union[bytes, SlabRef]:
        """What goes into a weights message: a descriptor on the fabric, else the reference pickle."""
        from myfyp_b200.settings import Settings

        if device_ok and Settings.DEVICE_WEIGHTS:
            return self.as_ref() if self.model is None else self.publish()
        return self.encode_parameters()

    def encode_parameters(self, params: Optional[List[np.ndarray]] = None) -> bytes:
        """The reference's pickle.  Slab-shaped device vectors inside ``additional_info`` (Scaffold's deltas and global c) leave
        in the reference's form - one array per ``state_dict`` entry (pytorch/callbacks/scaffold_callback.py:131-135) - so a
        reference node, and our own receiving side, read them layer by layer."""
        if params is None:
            params = self.get_parameters()
        return pickle.dumps({"params": params, "additional_info": _info_to_numpy(_slab_vectors_to_layers(self.additional_info, self.layout))})

    # ------------------------------------------------------------------ parameter access
    def get_parameters(self) -> List[np.ndarray]:
        self.sync_mirrors_to_slab()
        host = self.slab.detach().cpu().numpy()
        out = []
        for e in self.layout.in_order():
            a = host[e.offset : e.offset + e.numel].reshape(e.shape)
            if e.dtype != torch.float32:
                a = a.round().astype(_np_dtype(e.dtype)) if not e.dtype.is_floating_point else a.astype(_np_dtype(e.dtype))
            else:
                a = a.copy()
            out.append(a)
        return out

    def set_parameters(self, params: Union[List[np.ndarray], bytes, SlabRef, torch.Tensor, "TorchP2PFLModel"]) -> None:
        if isinstance(params, (bytes, bytearray, memoryview)):
            params, info = self.decode_parameters(bytes(params))
            self.additional_info.update(info)
        if isinstance(params, TorchP2PFLModel):
            if params.layout.hash != self.layout.hash:
                raise ModelNotMatchingError("Not matching models")
            self._copy_from_flat(params.slab)
        elif isinstance(params, SlabRef):
            if params.layout_hash != self.layout.hash or params.numel != self.layout.total:
                raise ModelNotMatchingError("Not matching models")
            self.additional_info.update(params.additional_info)
            self._copy_from_flat(params.resolve())
        elif isinstance(params, torch.Tensor):
            if params.numel() != self.layout.total:
                raise ModelNotMatchingError("Not matching models")
            self._copy_from_flat(params)
        else:
            entries = self.layout.in_order()
            params = list(params)
            if len(params) != len(entries):
                raise ModelNotMatchingError("Not matching models")
            host = np.zeros(self.layout.total, dtype=np.float32)
            for e, p in zip(entries, params):
                p = np.asarray(p)
                if tuple(p.shape) != tuple(e.shape):
                    raise ModelNotMatchingError("Not matching models")
                host[e.offset : e.offset + e.numel] = p.astype(np.float32, copy=False).reshape(-1)
            self._copy_from_flat(torch.from_numpy(host))
        self.sync_mirrors_from_slab()

    def _copy_from_flat(self, flat: torch.Tensor) -> None:
        from myfyp_b200 import ops

        ops.pull_copy(self.slab, flat)

    # ------------------------------------------------------------------ copies
    def build_copy(self, **kwargs) -> "TorchP2PFLModel":
        """A *held* model sharing this model's layout (no ``nn.Module`` deep copy)."""
        params = kwargs.pop("params", None)
        held = TorchP2PFLModel(None, layout=self.layout, device=self.slab.device, slab=_adoptable_slab(params, self), **kwargs)
        if params is not None and held.slab is not _as_tensor(params):
            held.set_parameters(params)
        elif isinstance(params, SlabRef):
            held.additional_info.update(params.additional_info)
            held._ref = params
        return held

    def snapshot_model(self) -> "TorchP2PFLModel":
        """Held, immutable copy of the live model (one D2D copy) with its contribution and info."""
        if self.model is None:
            return self
        return self.build_copy(params=self.publish(), num_samples=self.num_samples, contributors=list(self.contributors), additional_info=dict(self.additional_info))

    def get_framework(self) -> str:
        return "pytorch"


def _as_tensor(params: Any) -> Optional[torch.Tensor]:
    if isinstance(params, SlabRef):
        return params.tensor
    if isinstance(params, torch.Tensor):
        return params
    return None


def _adoptable_slab(params: Any, like: TorchP2PFLModel) -> Optional[torch.Tensor]:
    """A descriptor/tensor of the right size is adopted zero-copy (it is immutable by contract).

    It may live on another GPU: kernels dereference it over NVLink when they aggregate.
    """
    t = _as_tensor(params)
    if t is None:
        return None
    if isinstance(params, SlabRef) and params.layout_hash != like.layout.hash:
        raise ModelNotMatchingError("Not matching models")
    if t.numel() != like.layout.total or t.dtype != torch.float32:
        raise ModelNotMatchingError("Not matching models")
    return t


def _np_dtype(dt: torch.dtype):
    return {torch.int64: np.int64, torch.int32: np.int32, torch.float16: np.float16, torch.float64: np.float64, torch.uint8: np.uint8, torch.bool: np.bool_}.get(
        dt, np.float32
    )


# name used by scripts written against the reference
LightningModel = TorchP2PFLModel
