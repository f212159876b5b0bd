"""Turn a *local* torchvision dataset into a :class:`P2PFLDataset`.

The reference's utility converts MNIST / CIFAR / ... to a Hugging Face dataset and pushes it to the Hub
(learning/frameworks/pytorch/utils/torchvision_to_datasets.py:31-190).  There is no network here, so this reads a
torchvision dataset that is already on disk (``download=False``) and keeps it in memory; nothing is uploaded.

    python -m myfyp_b200.learning.dataset.torchvision_import --dataset CIFAR10 --root /data --out cifar10.npz
"""

from __future__ import annotations

import argparse
from typing import Optional

import numpy as np

from myfyp_b200.learning.dataset.p2pfl_dataset import P2PFLDataset

SUPPORTED = ["MNIST", "FashionMNIST", "EMNIST", "QMNIST", "CIFAR10", "CIFAR100"]


def from_torchvision(name: str, root: str, download: bool = False, limit: Optional[int] = None, **kwargs) -> P2PFLDataset:
    import torchvision

    if name not in SUPPORTED:
        raise ValueError(f"unsupported dataset {name}; choose from {SUPPORTED}")
    cls = getattr(torchvision.datasets, name)

    def split(train: bool):
        ds = cls(root=root, train=train, download=download, **kwargs)
        x = np.asarray(ds.data)[:limit]
        y = np.asarray(ds.targets)[:limit].astype(np.int64)
        x = x.astype(np.float32) / 255.0
        if x.ndim == 4:  # NHWC uint8 -> NCHW float
            x = x.transpose(0, 3, 1, 2)
        return {"image": x, "label": y}

    return P2PFLDataset.from_arrays(split(True), split(False))


def main() -> None:
    ap = argparse.ArgumentParser(description=__doc__.split("\n\n")[0])
    ap.add_argument("--dataset", required=True, choices=SUPPORTED)
    ap.add_argument("--root", required=True)
    ap.add_argument("--out", required=True, help=".npz file with train/test image and label arrays")
    a = ap.parse_args()
    d = from_torchvision(a.dataset, a.root)
    np.savez_compressed(a.out, **{f"{s}_{k}": d.get(slice(None), train=(s == "train"))[k] for s in ("train", "test") for k in ("image", "label")})
    print(f"wrote {a.out}")


if __name__ == "__main__":
    main()
