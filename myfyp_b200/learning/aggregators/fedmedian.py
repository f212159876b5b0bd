"""Coordinate-wise median / trimmed mean (Yin et al. 2018).

The reference's ``FedMedian.aggregate`` raises ``NotImplementedError`` as its first
statement (p2pfl/learning/aggregators/fedmedian.py:47); the dead code below it is
``np.median(np.array(layer_weights), axis=0)`` per layer (:55-65) and forgets the
contributors.  This is the working version: one K-way kernel (A5) where every thread
holds the K values of a coordinate in registers and runs a sorting network.

``partial_aggregation=False``: a median of partial means is not the median (the
reference's docs table claims otherwise).
"""

from __future__ import annotations

from typing import List

from myfyp_b200.learning.aggregators.aggregator import Aggregator, NoModelsToAggregateError
from myfyp_b200.learning.p2pfl_model import P2PFLModel


class FedMedian(Aggregator):
    """``trim=None`` -> exact coordinate-wise median; ``trim=b`` -> mean after dropping the b smallest and b largest."""

    def __init__(self, node_name: str = "unknown", trim: int | None = None) -> None:
        super().__init__(node_name)
        self.partial_aggregation = False
        self.trim = trim

    def aggregate(self, models: List[P2PFLModel]) -> P2PFLModel:
        if len(models) == 0:
            raise NoModelsToAggregateError(f"({self.node_name}) Trying to aggregate models when there is no models")
        from myfyp_b200 import ops
        from myfyp_b200.learning.p2pfl_model import TorchP2PFLModel

        total = sum(m.get_num_samples() for m in models)
        contributors: List[str] = []
        for m in models:
            contributors += m.get_contributors()
        trim = self.trim
        if trim is not None and 2 * trim >= len(models):
            trim = (len(models) - 1) // 2
        if all(isinstance(m, TorchP2PFLModel) for m in models):
            first: TorchP2PFLModel = models[0]  # type: ignore[assignment]
            slabs = [m.slab for m in models]  # type: ignore[attr-defined]
            out = ops.median(slabs, trim=trim, device=ops.pick_device(slabs, prefer=self.local_device(models)))
            return first.build_copy(params=out, num_samples=total, contributors=contributors)
        import numpy as np

        layers = [m.get_parameters() for m in models]
        out_layers = []
        for i in range(len(layers[0])):
            stack = np.sort(np.stack([l[i] for l in layers]).astype(np.float64), axis=0)
            red = np.median(stack, axis=0) if trim is None else stack[trim : len(models) - trim].mean(axis=0)
            out_layers.append(red.astype(layers[0][i].dtype))
        return models[0].build_copy(params=out_layers, num_samples=total, contributors=contributors)


class TrimmedMean(FedMedian):
    def __init__(self, node_name: str = "unknown", trim: int = 1) -> None:
        super().__init__(node_name, trim=trim)
