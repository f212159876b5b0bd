"""A small CIFAR-10 CNN (BASELINE.json config 2).  The reference ships no CNN (only the thesis describes a
FEMNIST one), so the architecture is defined here: 3x[conv3x3-ReLU-maxpool2] + 2 linear layers, ~0.55 M
parameters (about the "9x the MLP"... order of magnitude the thesis quotes for its CNN)."""

from __future__ import annotations

from typing import Optional

import torch
from torch import nn


class CNN(nn.Module):
    def __init__(self, in_channels: int = 3, num_classes: int = 10, width: int = 32, image_size: int = 32) -> None:
        super().__init__()
        w = width
        self.features = nn.Sequential(
            nn.Conv2d(in_channels, w, 3, padding=1), nn.ReLU(inplace=True), nn.MaxPool2d(2),
            nn.Conv2d(w, 2 * w, 3, padding=1), nn.ReLU(inplace=True), nn.MaxPool2d(2),
            nn.Conv2d(2 * w, 4 * w, 3, padding=1), nn.ReLU(inplace=True), nn.MaxPool2d(2),
        )
        s = image_size // 8
        self.classifier = nn.Sequential(nn.Flatten(), nn.Linear(4 * w * s * s, 256), nn.ReLU(inplace=True), nn.Linear(256, num_classes))

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return self.classifier(self.features(x.float()))


def cnn_model(device: Optional[str] = None, seed: Optional[int] = None, native: Optional[bool] = None, precision: str = "tf32", **kwargs):
    """``TorchP2PFLModel`` around a fresh CNN; ``native`` (default: on CUDA) attaches a hand-written training step which the
    learner prefers over autograd / cuDNN:

    * ``precision="tf32"`` (default): fp32 storage, TF32 tensor-core convolutions, fp32-class linear layers - torch's own
      defaults for this module - trained node-batched (:mod:`myfyp_b200.models.grouped_cnn`, ``csrc/train_tf32.cu``);
    * ``precision="bf16"``: bf16 operands / fp32 accumulation, one graph per node (:mod:`myfyp_b200.models.native_cnn`)."""
    from myfyp_b200.learning.p2pfl_model import TorchP2PFLModel

    if seed is not None:
        torch.manual_seed(seed)
    module = CNN(**kwargs)
    module.image_size = kwargs.get("image_size", 32)
    model = TorchP2PFLModel(module, device=device)
    if native is None:
        module.wants_native = True
        native = model.device.type == "cuda"
    if native:
        spec = None
        if precision == "tf32":
            from myfyp_b200.models.grouped_cnn import attach_grouped_cnn

            spec = attach_grouped_cnn(model, module.image_size)
        elif precision != "bf16":
            raise ValueError("precision must be 'tf32' or 'bf16'")
        if spec is None:  # bf16 asked for, or an architecture outside the node-batched plan
            from myfyp_b200.models.native_cnn import attach_native_cnn

            attach_native_cnn(model, module.image_size)
        module.precision = "tf32" if spec is not None else "bf16"
    return model
