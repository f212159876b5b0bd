"""The reference's example MLP: ``Linear(784,256)-ReLU-Linear(256,128)-ReLU-Linear(128,10)``
(p2pfl/learning/frameworks/pytorch/lightning_model.py:145-179), 235 146 parameters in 6 tensors.

``MLP`` is a plain ``nn.Module`` (autograd path).  On a CUDA device :func:`mlp_model` returns it with the
hand-written tcgen05 training step attached (``native_train_step`` / ``native_eval``), which the learner
prefers over autograd (see :mod:`myfyp_b200.models.native_mlp`).
"""

from __future__ import annotations

from typing import Optional, Sequence

import torch
from torch import nn


class MLP(nn.Module):
    def __init__(self, input_size: int = 28 * 28, hidden_sizes: Sequence[int] = (256, 128), out_channels: int = 10) -> None:
        super().__init__()
        sizes = [input_size, *hidden_sizes, out_channels]
        self.layers = nn.ModuleList(nn.Linear(a, b) for a, b in zip(sizes[:-1], sizes[1:]))
        self.sizes = sizes

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x = x.reshape(x.shape[0], -1).float()
        for i, layer in enumerate(self.layers):
            x = layer(x)
            if i + 1 < len(self.layers):
                x = torch.relu(x)
        # the reference applies log_softmax here and cross_entropy on top (double log-softmax, idempotent,
        # SURVEY Q12); we return logits and fuse log-softmax + NLL in the loss kernel
        return x


def mlp_model(device: Optional[str] = None, seed: Optional[int] = None, native: Optional[bool] = None, **kwargs):
    """``TorchP2PFLModel`` around a fresh MLP; ``native`` (default: on CUDA) attaches the tcgen05 training step."""
    from myfyp_b200.learning.p2pfl_model import TorchP2PFLModel

    if seed is not None:
        torch.manual_seed(seed)
    module = MLP(**kwargs)
    model = TorchP2PFLModel(module, device=device)
    if native is None:
        module.wants_native = True  # attach the tcgen05 path whenever the model lands on a GPU (Node places it)
        native = model.device.type == "cuda"
    if native:
        from myfyp_b200.models.native_mlp import attach_native_mlp

        attach_native_mlp(model)
    return model
