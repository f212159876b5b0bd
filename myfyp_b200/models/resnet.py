"""ResNet-18, CIFAR variant (3x3 stem, no max-pool), 11.17 M parameters (BASELINE.json configs 3-4).
Not in the reference.  BatchNorm running statistics and ``num_batches_tracked`` travel in the
slab like every other ``state_dict`` entry (the reference averages them as floats)."""

from __future__ import annotations

from typing import Optional

import torch
from torch import nn


class BasicBlock(nn.Module):
    def __init__(self, cin: int, cout: int, stride: int) -> None:
        super().__init__()
        self.conv1 = nn.Conv2d(cin, cout, 3, stride, 1, bias=False)
        self.bn1 = nn.BatchNorm2d(cout)
        self.conv2 = nn.Conv2d(cout, cout, 3, 1, 1, bias=False)
        self.bn2 = nn.BatchNorm2d(cout)
        self.shortcut = None
        if stride != 1 or cin != cout:
            self.shortcut = nn.Sequential(nn.Conv2d(cin, cout, 1, stride, bias=False), nn.BatchNorm2d(cout))

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        out = torch.relu(self.bn1(self.conv1(x)))
        out = self.bn2(self.conv2(out))
        return torch.relu(out + (x if self.shortcut is None else self.shortcut(x)))


class ResNet18(nn.Module):
    def __init__(self, num_classes: int = 10, in_channels: int = 3, width: int = 64) -> None:
        super().__init__()
        self.conv1 = nn.Conv2d(in_channels, width, 3, 1, 1, bias=False)
        self.bn1 = nn.BatchNorm2d(width)
        layers, cin = [], width
        for i, (mult, stride) in enumerate([(1, 1), (2, 2), (4, 2), (8, 2)]):
            cout = width * mult
            layers += [BasicBlock(cin, cout, stride), BasicBlock(cout, cout, 1)]
            cin = cout
        self.layers = nn.Sequential(*layers)
        self.fc = nn.Linear(cin, num_classes)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x = torch.relu(self.bn1(self.conv1(x.float())))
        x = self.layers(x)
        return self.fc(torch.flatten(torch.nn.functional.adaptive_avg_pool2d(x, 1), 1))


def resnet18_model(device: Optional[str] = None, seed: Optional[int] = None, native: Optional[bool] = None, image_size: int = 32, **kwargs):
    """``TorchP2PFLModel`` around a fresh ResNet-18; ``native`` (default: on CUDA) attaches the hand-written training step
    (:mod:`myfyp_b200.models.native_resnet`) which the learner prefers over autograd / cuDNN."""
    from myfyp_b200.learning.p2pfl_model import TorchP2PFLModel

    if seed is not None:
        torch.manual_seed(seed)
    module = ResNet18(**kwargs)
    module.image_size = image_size
    model = TorchP2PFLModel(module, device=device)
    if native is None:
        module.wants_native = True
        native = model.device.type == "cuda"
    if native:
        from myfyp_b200.models.native_resnet import attach_native_resnet

        attach_native_resnet(model, image_size)
    return model
