"""SCAFFOLD client side.

Reference: pytorch/callbacks/scaffold_callback.py:54-140 (whose gradient correction is a
no-op because it edits ``state_dict()`` copies) and the Keras twin
tensorflow/callbacks/scaffold_callback.py:49-59 (which does apply it).

Per fit:  start: ``c <- additional_info["global_c"]`` (else keep / zeros), ``x <- theta``, ``K <- 0``;
every step the fused optimizer adds the control-variate term to the gradient (no extra
kernel, T6); end (kernel T7): ``c_i <- c_i + (x - y)/(K*lr)``, ``delta_y = y - x``, ``delta_c = c_i_new - c_i_old``.

``variant="paper"`` (default): ``g += c - c_i`` (Karimireddy et al., Alg. 1).
``variant="reference"``: ``g += lr * (c_i - c)`` - what the reference's Keras wrapper computes.
"""

from __future__ import annotations

from typing import Any, Dict

import numpy as np
import torch

from myfyp_b200.learning.callbacks.callback import P2PFLCallback


class SCAFFOLDCallback(P2PFLCallback):
    def __init__(self, aggregator=None, variant: str = "paper") -> None:
        super().__init__()
        self.variant = variant
        self.c_i: torch.Tensor | None = None
        self.c: torch.Tensor | None = None
        self.x_start: torch.Tensor | None = None

    @staticmethod
    def get_name() -> str:
        return "scaffold"

    def on_fit_start(self, learner) -> None:
        model = learner.get_model()
        n = model.layout.n_trainable
        slab = model.slab
        if self.c_i is None or self.c_i.numel() != n or self.c_i.device != slab.device:
            self.c_i = torch.zeros(n, dtype=torch.float32, device=slab.device)
        # c and x_start are allocated once and refreshed in place: captured CUDA graphs keep pointing at them
        if self.c is None or self.c.numel() != n or self.c.device != slab.device:
            self.c = torch.zeros(n, dtype=torch.float32, device=slab.device)
        g = self.additional_info.get("global_c") if isinstance(self.additional_info, dict) else None
        if g is not None:
            self.c.copy_(_flat_prefix(g, model, n))
        # x_start covers the whole slab: the non-trainable tail (BatchNorm running statistics) is part of delta_y too,
        # otherwise the aggregated model keeps the buffers it was initialised with
        if self.x_start is None or self.x_start.numel() != slab.numel() or self.x_start.device != slab.device:
            self.x_start = slab.clone()
        else:
            self.x_start.copy_(slab)

    def optimizer_terms(self, learner) -> Dict[str, Any]:
        if self.variant == "reference":
            return {"cv_global": self.c, "cv_local": self.c_i, "cv_scale": -float(learner.lr)}
        return {"cv_global": self.c, "cv_local": self.c_i, "cv_scale": 1.0}

    def on_fit_end(self, learner, steps: int, lr: float) -> None:
        from myfyp_b200 import ops

        model = learner.get_model()
        n = model.layout.n_trainable
        total = model.layout.total
        dy = torch.zeros(total, dtype=torch.float32, device=model.slab.device)
        dc = torch.zeros(total, dtype=torch.float32, device=model.slab.device)
        ops.scaffold_finalize(self.c_i, dy[:n], dc[:n], self.x_start[:n], model.slab[:n], max(steps, 1) * lr)
        if total > n:
            torch.sub(model.slab[n:], self.x_start[n:], out=dy[n:])
        # flat device slabs on the fabric; `encode_parameters` turns them into NumPy for the wire
        self.additional_info = {"delta_y_i": dy, "delta_c_i": dc}


def _flat_prefix(value: Any, model, n: int) -> torch.Tensor:
    from myfyp_b200.parallel.slab import flat_from_wire

    return flat_from_wire(value, model.layout, model.slab.device)[:n].contiguous()
