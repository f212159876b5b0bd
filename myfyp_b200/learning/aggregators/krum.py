"""Krum / Multi-Krum (Blanchard et al. 2017).  Not in the reference.

``score_i = sum of the K-f-2 smallest ||theta_i - theta_j||^2``; Krum adopts the argmin,
Multi-Krum averages the ``m`` best (sample-count weighted, like FedAvg).

Device path: pairwise distances come from a Gram matrix ``G = D D^T`` of the centred
models ``D_k = theta_k - theta_0`` computed on tcgen05 tensor cores (kernel A6: split-K
over coordinate chunks, TMEM accumulators), ``d_ij = G_ii + G_jj - 2 G_ij``; scoring and
selection are one tiny kernel (A7); the selected subset is averaged with the FedAvg
kernel.  Centring keeps the Gram trick away from catastrophic cancellation when the
models are close (late rounds).
"""

from __future__ import annotations

from typing import List

from myfyp_b200.learning.aggregators.aggregator import Aggregator, NoModelsToAggregateError
from myfyp_b200.learning.aggregators.fedavg import weighted_average
from myfyp_b200.learning.p2pfl_model import P2PFLModel


class Krum(Aggregator):
    def __init__(self, node_name: str = "unknown", f: int = 0, m: int = 1) -> None:
        super().__init__(node_name)
        self.partial_aggregation = False
        self.f = f
        self.m = m
        self.last_scores: List[float] = []
        self.last_selected: List[int] = []

    def aggregate(self, models: List[P2PFLModel]) -> P2PFLModel:
        if len(models) == 0:
            raise NoModelsToAggregateError(f"({self.node_name}) Trying to aggregate models when there is no models")
        from myfyp_b200 import ops
        from myfyp_b200.learning.p2pfl_model import TorchP2PFLModel

        k = len(models)
        m = max(1, min(self.m, k))
        if all(isinstance(x, TorchP2PFLModel) for x in models):
            slabs = [x.slab for x in models]  # type: ignore[attr-defined]
            dist = ops.pairwise_sqdist(slabs, device=ops.pick_device(slabs, prefer=self.local_device(models)))
            scores, selected = ops.krum_select(dist, self.f, m)
        else:
            import numpy as np

            flat = [np.concatenate([np.asarray(l, dtype=np.float64).ravel() for l in x.get_parameters()]) for x in models]
            dist_np = np.array([[float(((a - b) ** 2).sum()) for b in flat] for a in flat])
            scores, selected = krum_select_reference(dist_np, self.f, m)
        self.last_scores, self.last_selected = list(scores), list(selected)
        chosen = [models[i] for i in selected]
        total = sum(x.get_num_samples() for x in chosen)
        # the result stands for the whole train set (everyone was considered), so the aggregate
        # carries every contributor - otherwise peers would keep waiting for the rejected ones
        contributors: List[str] = []
        for x in models:
            contributors += x.get_contributors()
        all_samples = sum(x.get_num_samples() for x in models)
        return weighted_average(chosen, [x.get_num_samples() / total for x in chosen], all_samples, contributors, device=self.local_device(models))


def krum_select_reference(dist, f: int, m: int):
    """NumPy oracle: scores and the indices of the ``m`` best (ties -> lower index)."""
    import numpy as np

    dist = np.asarray(dist, dtype=np.float64)
    dist = np.where(np.isnan(dist), np.inf, dist)  # a broken model is infinitely far away
    k = dist.shape[0]
    closest = max(1, k - f - 2)
    scores = []
    for i in range(k):
        others = np.sort(np.delete(dist[i], i))
        finite = int(np.isfinite(others).sum())
        take = min(closest, k - 1, finite) if finite else min(1, k - 1)
        scores.append(float(others[:take].sum()) if k > 1 else 0.0)
    order = sorted(range(k), key=lambda i: (scores[i], i))
    return scores, order[:m]
