"""Sample-count-weighted Federated Averaging (McMahan et al. 2016).

Reference math: p2pfl/learning/aggregators/fedavg.py:41-77 -
``out = sum_k n_k * theta_k / sum_k n_k`` per layer with NumPy on the host.
Here it is one K-way pull-reduce kernel over the flat slabs (kernel A1): sources may be
local or peer-GPU slabs, the weights ``n_k / sum n`` are kernel arguments, output is a
held model.  A single source is returned as is (no copy).  ``partial_aggregation=True``.
"""

from __future__ import annotations

from typing import List

from myfyp_b200.learning.aggregators.aggregator import Aggregator, NoModelsToAggregateError
from myfyp_b200.learning.p2pfl_model import P2PFLModel


class FedAvg(Aggregator):
    def __init__(self, node_name: str = "unknown") -> None:
        super().__init__(node_name)
        self.partial_aggregation = True

    def aggregate(self, models: List[P2PFLModel]) -> P2PFLModel:
        if len(models) == 0:
            raise NoModelsToAggregateError(f"({self.node_name}) Trying to aggregate models when there is no models")
        total = sum(m.get_num_samples() for m in models)
        contributors: List[str] = []
        for m in models:
            contributors += m.get_contributors()
        return weighted_average(models, [m.get_num_samples() / total for m in models], total, contributors, device=self.local_device(models))


def weighted_average(models: List[P2PFLModel], weights: List[float], num_samples: int, contributors: List[str], device=None) -> P2PFLModel:
    """``sum_k w_k theta_k`` as a held model; device kernel for slab models (on ``device``, default: the first model's), NumPy otherwise."""
    from myfyp_b200 import ops
    from myfyp_b200.learning.p2pfl_model import TorchP2PFLModel

    if all(isinstance(m, TorchP2PFLModel) for m in models):
        first: TorchP2PFLModel = models[0]  # type: ignore[assignment]
        if len(models) == 1 and abs(weights[0] - 1.0) < 1e-12:
            # one model, weight 1: hand out the same immutable slab (with its arena coordinates, if any) - no kernel, no copy
            return first.build_copy(params=first.as_ref() if first.model is None else first.publish(), num_samples=num_samples, contributors=contributors)
        dev = ops.pick_device([m.slab for m in models], prefer=device if device is not None else first.slab.device)  # type: ignore[attr-defined]
        out = ops.fedavg([m.slab for m in models], weights, device=dev)  # type: ignore[attr-defined]
        like = next((m for m in models if m.slab.device == dev), first)  # type: ignore[attr-defined]
        return like.build_copy(params=out, num_samples=num_samples, contributors=contributors)
    # generic (non-torch) models: per-layer NumPy, like the reference
    import numpy as np

    acc = [np.zeros_like(layer, dtype=np.float64) for layer in models[0].get_parameters()]
    for m, w in zip(models, weights):
        for i, layer in enumerate(m.get_parameters()):
            acc[i] += np.asarray(layer, dtype=np.float64) * w
    ref = models[0].get_parameters()
    return models[0].build_copy(params=[a.astype(r.dtype) for a, r in zip(acc, ref)], num_samples=num_samples, contributors=contributors)
