"""SCAFFOLD aggregator (Karimireddy et al. 2019).

Reference math, replicated exactly because its tests pin it
(p2pfl/learning/aggregators/scaffold.py:70-122, test/learning/scaffold_test.py:80-141):

    x <- x + lr_g * sum_k (n_k / sum n) * delta_y_k          (x starts as models[0]'s parameters)
    c <- c + (1/K) * sum_k delta_c_k                          (c starts at 0)

and the result carries ``additional_info["scaffold"] = {"global_c": c}``.
``delta_y_i`` / ``delta_c_i`` arrive in each model's ``additional_info["scaffold"]``, either as
lists of per-layer arrays (wire form) or as flat device slabs (fabric form); device form runs as one
fused K-way kernel (A8) over both sets of slabs.  ``partial_aggregation=False``.
"""

from __future__ import annotations

from typing import Any, Dict, List

import numpy as np
import torch

from myfyp_b200.learning.aggregators.aggregator import Aggregator, NoModelsToAggregateError
from myfyp_b200.learning.p2pfl_model import P2PFLModel


class Scaffold(Aggregator):
    REQUIRED_INFO_KEYS = ["delta_y_i", "delta_c_i"]

    def __init__(self, node_name: str = "unknown", global_lr: float = 0.1, rebase: bool = True, variate_scaling: str = "reference") -> None:
        """``variate_scaling``: how the clients' ``delta_c`` enter the global control variate.

        * ``"reference"`` (default): ``c += (1/K) sum_k delta_c_k`` over the K contributors of the round - the reference's formula
          (scaffold.py:90-111, pinned by its test).  With every node in every train set this is the paper's; with a train set smaller
          than the federation it adds the participants' mean at full weight every round, ``c`` stops being the average of the ``c_i``
          and grows round after round - with 8 of 64 nodes per round training falls to chance within a few rounds;
        * ``"paper"``: ``c += (1/N) sum_k delta_c_k`` with N = the size of the federation (Karimireddy et al., Alg. 1, line 17), which
          keeps ``c = mean_i c_i``.  N is ``federation_size`` (set by the train stage from the neighbour table every round)."""
        super().__init__(node_name)
        if variate_scaling not in ("reference", "paper"):
            raise ValueError("variate_scaling must be 'reference' or 'paper'")
        self.variate_scaling = variate_scaling
        self.global_lr = global_lr
        self.rebase = rebase  # slab path: recover the round's starting point from the clients' models (see _aggregate_slab)
        self.c: Any = []  # list of arrays (generic path) or flat tensor (slab path)
        self.global_model_params: Any = []
        self.partial_aggregation = False

    def get_required_callbacks(self) -> List[str]:
        return ["scaffold"]

    def _variate_weight(self, k: int) -> float:
        if self.variate_scaling == "paper":
            return 1.0 / max(int(getattr(self, "federation_size", None) or k), k)
        return 1.0 / k

    def observe_global(self, model: P2PFLModel) -> None:
        try:
            g = model.get_info("scaffold").get("global_c")
        except Exception:  # noqa: BLE001
            return
        if isinstance(g, torch.Tensor):
            self.c = g.detach().clone()
        elif g is not None and hasattr(model, "layout"):
            self.c = _as_flat(g, model)

    def _get_and_validate_model_info(self, model: P2PFLModel) -> Dict[str, Any]:
        info = model.get_info("scaffold")
        if not all(k in info for k in self.REQUIRED_INFO_KEYS):
            raise ValueError(f"Model is missing required info keys: {self.REQUIRED_INFO_KEYS}. Model info keys: {list(info.keys())}")
        return info

    def aggregate(self, models: List[P2PFLModel]) -> P2PFLModel:
        if not models:
            raise NoModelsToAggregateError(f"({self.node_name}) Trying to aggregate models when there is no models")
        from myfyp_b200.learning.p2pfl_model import TorchP2PFLModel

        infos = [self._get_and_validate_model_info(m) for m in models]
        total = sum(m.get_num_samples() for m in models)
        contributors: List[str] = []
        for m in models:
            contributors.extend(m.get_contributors())
        if all(isinstance(m, TorchP2PFLModel) for m in models):
            return self._aggregate_slab(models, infos, total, contributors)  # type: ignore[arg-type]
        return self._aggregate_numpy(models, infos, total, contributors)

    # ------------------------------------------------------------------ generic path (reference math in NumPy)
    def _aggregate_numpy(self, models, infos, total, contributors) -> P2PFLModel:
        first = models[0].get_parameters()
        acc_y = [np.zeros_like(l, dtype=np.float64) for l in first]
        for m, info in zip(models, infos):
            for i, layer in enumerate(info["delta_y_i"]):
                acc_y[i] += np.asarray(layer, dtype=np.float64) * m.get_num_samples()
        acc_y = [(l / total) * self.global_lr for l in acc_y]
        if not isinstance(self.global_model_params, list) or not self.global_model_params:
            self.global_model_params = [np.asarray(p) for p in first]
        self.global_model_params = [(p + d).astype(p.dtype) for p, d in zip(self.global_model_params, acc_y)]
        acc_c = [np.zeros_like(np.asarray(l), dtype=np.float64) for l in infos[0]["delta_c_i"]]
        for info in infos:
            for i, layer in enumerate(info["delta_c_i"]):
                acc_c[i] += np.asarray(layer, dtype=np.float64)
        acc_c = [l * self._variate_weight(len(models)) for l in acc_c]
        if not isinstance(self.c, list) or not self.c:
            self.c = [np.zeros_like(l) for l in acc_c]
        self.c = [c + d for c, d in zip(self.c, acc_c)]
        out = models[0].build_copy(params=self.global_model_params, num_samples=total, contributors=contributors)
        out.add_info("scaffold", {"global_c": self.c})
        return out

    # ------------------------------------------------------------------ slab path (kernel A8)
    def _aggregate_slab(self, models, infos, total, contributors) -> P2PFLModel:
        from myfyp_b200 import ops

        first = models[0]
        dev = self.local_device(models)
        dys = [_as_flat(info["delta_y_i"], first) for info in infos]
        dcs = [_as_flat(info["delta_c_i"], first) for info in infos]
        ops.order_after_producers([first.slab, *dys, *dcs], dev)  # clients' tensors may live on other GPUs of this process
        if self.rebase:
            # x = the model every client started this round from = y_k - delta_y_k (any k).  A copy kept from "our" last
            # aggregation is stale on a node that was outside the train set for a few rounds (train set < network).
            self.global_model_params = first.slab.to(dev, copy=True) - dys[0].to(dev)
        elif not isinstance(self.global_model_params, torch.Tensor):
            self.global_model_params = first.slab.to(dev, copy=True)
        if not isinstance(self.c, torch.Tensor) or self.c.numel() != self.global_model_params.numel():
            self.c = torch.zeros_like(self.global_model_params)
        elif self.c.device != dev:
            self.c = self.c.to(dev)
        wy = [self.global_lr * m.get_num_samples() / total for m in models]
        ops.scaffold_server_update(self.global_model_params, self.c, dys, dcs, wy, self._variate_weight(len(models)))
        out = first.build_copy(params=self.global_model_params.clone(), num_samples=total, contributors=contributors)
        out.add_info("scaffold", {"global_c": self.c.clone()})
        return out


def _as_flat(value: Any, like) -> torch.Tensor:
    """delta in wire form (list of per-layer arrays), flat form (1-D array of slab length) or fabric form (flat tensor) -> flat
    tensor on the model's device; a shape that is none of these raises."""
    from myfyp_b200.parallel.slab import flat_from_wire

    if isinstance(value, torch.Tensor) and value.numel() == like.layout.total:
        return value
    return flat_from_wire(value, like.layout, like.slab.device)
