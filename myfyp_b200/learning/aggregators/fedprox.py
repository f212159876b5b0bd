"""FedProx (Li et al. 2018).  Not in the reference.

Server side FedProx *is* FedAvg; the method lives in the local objective
``loss + (mu/2) * ||theta - theta_global||^2``, i.e. ``g += mu * (theta - theta_global)``.
The aggregator therefore only asks the learner for the ``"fedprox"`` callback, which
snapshots ``theta_global`` at the start of every fit and feeds it, with ``mu``, to the fused
optimizer kernel (T6) - the proximal term costs no extra kernel.
"""

from __future__ import annotations

from typing import List

from myfyp_b200.learning.aggregators.fedavg import FedAvg


class FedProx(FedAvg):
    def __init__(self, node_name: str = "unknown", mu: float = 0.01) -> None:
        super().__init__(node_name)
        self.mu = mu

    def get_required_callbacks(self) -> List[str]:
        return ["fedprox"]
