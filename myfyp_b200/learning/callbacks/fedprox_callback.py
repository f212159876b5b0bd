"""FedProx client side: ``g += mu * (theta - theta_global)`` inside the fused optimizer (T6).

Not in the reference.  ``theta_global`` is the slab at the start of the fit.
"""

from __future__ import annotations

from typing import Any, Dict

import torch

from myfyp_b200.learning.callbacks.callback import P2PFLCallback


class FedProxCallback(P2PFLCallback):
    def __init__(self, aggregator=None, mu: float | None = None) -> None:
        super().__init__()
        self.mu = float(mu if mu is not None else getattr(aggregator, "mu", 0.01))
        self.ref: torch.Tensor | None = None

    @staticmethod
    def get_name() -> str:
        return "fedprox"

    def on_fit_start(self, learner) -> None:
        model = learner.get_model()
        cur = model.slab[: model.layout.n_trainable]
        if self.ref is None or self.ref.shape != cur.shape or self.ref.device != cur.device:
            self.ref = cur.clone()  # allocated once: captured CUDA graphs keep pointing at it
        else:
            self.ref.copy_(cur)

    def optimizer_terms(self, learner) -> Dict[str, Any]:
        return {"prox_mu": self.mu, "prox_ref": self.ref}
