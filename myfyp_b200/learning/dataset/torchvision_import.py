"""torchvision dataset -> :class:`P2PFLDataset` / parquet shards / Hugging Face Hub.

Role of the reference's ``torchvision_to_datasets.py`` (p2pfl/learning/frameworks/pytorch/utils/torchvision_to_datasets.py:31-190: build
a Hugging Face ``DatasetDict`` from MNIST / CIFAR / ..., write a dataset card, push to the Hub).  Three layers here:

* :func:`from_torchvision` - in-memory :class:`P2PFLDataset` (what the examples use);
* :func:`export_parquet`   - ``train.parquet`` / ``test.parquet`` (columns ``image`` = PNG bytes in the Hub's ``Image`` struct, ``label``)
  plus a ``README.md`` dataset card with the class names; :meth:`P2PFLDataset.from_parquet` and ``datasets.load_dataset("parquet")``
  read them back;
* :func:`push_to_hub`      - uploads that folder as a dataset repository (needs ``huggingface_hub``, a token and a network; this image
  has no network, so the CLI says so instead of failing half-way).

CLI (same flags as the reference's script, plus ``--out_dir`` / ``--no_download`` / ``--limit``)::

    python -m myfyp_b200.learning.dataset.torchvision_import --dataset_name CIFAR10 --cache_dir ./data --out_dir ./cifar10_hf \\
        [--token hf_... --repo_id user/cifar10 --public] [--dataset_summary "..."] [--dataset_description "..."] [--official_link URL]
"""

from __future__ import annotations

import argparse
import io
import os
from typing import Any, Dict, List, Optional

import numpy as np

from myfyp_b200.learning.dataset.p2pfl_dataset import P2PFLDataset

SUPPORTED = ["MNIST", "FashionMNIST", "EMNIST", "QMNIST", "CIFAR10", "CIFAR100"]


def _load_split(name: str, root: str, train: bool, download: bool, **kwargs):
    import torchvision

    if name not in SUPPORTED:
        raise ValueError(f"unsupported dataset {name}; choose from {SUPPORTED}")
    cls = getattr(torchvision.datasets, name)
    if name == "EMNIST":
        kwargs.setdefault("split", "balanced")
    if name == "QMNIST":  # its constructor has no ``train`` keyword
        return cls(root=root, what="train" if train else "test", download=download, **kwargs)
    return cls(root=root, train=train, download=download, **kwargs)


def _arrays(ds, limit: Optional[int]) -> Dict[str, np.ndarray]:
    """uint8 images as stored (NHW or NHWC) and int64 labels."""
    x = np.asarray(ds.data)[:limit]
    y = np.asarray(ds.targets)[:limit]
    if y.ndim > 1:  # QMNIST keeps extended labels
        y = y[:, 0]
    return {"image": x, "label": y.astype(np.int64)}


def from_torchvision(name: str, root: str, download: bool = False, limit: Optional[int] = None, **kwargs) -> P2PFLDataset:
    """Float32 NCHW (or NHW) images scaled to [0, 1] + int64 labels, train and test splits, all in memory."""

    def split(train: bool):
        a = _arrays(_load_split(name, root, train, download, **kwargs), limit)
        x = a["image"].astype(np.float32) / 255.0
        if x.ndim == 4:  # NHWC uint8 -> NCHW float
            x = x.transpose(0, 3, 1, 2)
        return {"image": x, "label": a["label"]}

    return P2PFLDataset.from_arrays(split(True), split(False))


# ---------------------------------------------------------------------------------------------------------------- parquet + card
def _png(img: np.ndarray) -> bytes:
    from PIL import Image

    buf = io.BytesIO()
    Image.fromarray(img).save(buf, format="PNG")
    return buf.getvalue()


def write_parquet_split(arrays: Dict[str, np.ndarray], path: str, class_names: Optional[List[str]] = None, shard_rows: int = 8192) -> int:
    """One parquet file in the layout the Hub's viewer and ``datasets`` understand: ``image: struct<bytes, path>``, ``label: int64``."""
    import json

    import pyarrow as pa
    import pyarrow.parquet as pq

    image_t = pa.struct([("bytes", pa.binary()), ("path", pa.string())])
    schema = pa.schema([("image", image_t), ("label", pa.int64())])
    features: Dict[str, Any] = {"image": {"_type": "Image"}, "label": {"_type": "ClassLabel", "names": class_names} if class_names else {"dtype": "int64", "_type": "Value"}}
    schema = schema.with_metadata({"huggingface": json.dumps({"info": {"features": features}})})
    n = len(arrays["label"])
    with pq.ParquetWriter(path, schema, compression="zstd") as w:
        for s in range(0, n, shard_rows):
            imgs = [{"bytes": _png(arrays["image"][i]), "path": None} for i in range(s, min(s + shard_rows, n))]
            w.write_table(pa.table({"image": pa.array(imgs, type=image_t), "label": pa.array(arrays["label"][s : s + shard_rows], type=pa.int64())}, schema=schema))
    return n


def dataset_card(name: str, n_train: int, n_test: int, class_names: Optional[List[str]], summary: str = "", description: str = "", link: str = "") -> str:
    head = ["---", "task_categories:", "- image-classification", f"pretty_name: {name}", "configs:", "- config_name: default", "  data_files:",
            "  - split: train", "    path: train.parquet", "  - split: test", "    path: test.parquet", "---", ""]
    body = [f"# {name}", "", summary or f"{name} converted from torchvision by myfyp_b200 (`learning/dataset/torchvision_import.py`).", ""]
    if description:
        body += [description, ""]
    body += [f"* train: {n_train} examples", f"* test: {n_test} examples"]
    if class_names:
        body.append("* classes: " + ", ".join(str(c) for c in class_names))
    if link:
        body += ["", f"Source: {link}"]
    body += ["", "Load it with `P2PFLDataset.from_parquet({'train': 'train.parquet', 'test': 'test.parquet'})` or",
             "`datasets.load_dataset('parquet', data_files=...)`.", ""]
    return "\n".join(head + body)


def export_parquet(name: str, root: str, out_dir: str, download: bool = False, limit: Optional[int] = None, summary: str = "", description: str = "", link: str = "", **kwargs) -> str:
    """``out_dir/{train,test}.parquet`` + ``README.md``; returns ``out_dir``."""
    os.makedirs(out_dir, exist_ok=True)
    counts, names = {}, None
    for split, train in (("train", True), ("test", False)):
        ds = _load_split(name, root, train, download, **kwargs)
        names = names or ([str(c) for c in ds.classes] if getattr(ds, "classes", None) is not None else None)
        counts[split] = write_parquet_split(_arrays(ds, limit), os.path.join(out_dir, f"{split}.parquet"), names)
    with open(os.path.join(out_dir, "README.md"), "w") as f:
        f.write(dataset_card(name, counts["train"], counts["test"], names, summary, description, link))
    return out_dir


def push_to_hub(folder: str, repo_id: str, token: str, public: bool = False) -> str:
    """Create (or reuse) the dataset repository ``repo_id`` and upload ``folder``.  Raises with a plain message when the Hub cannot
    be reached."""
    from huggingface_hub import HfApi

    api = HfApi(token=token)
    try:
        api.create_repo(repo_id, repo_type="dataset", private=not public, exist_ok=True)
        api.upload_folder(folder_path=folder, repo_id=repo_id, repo_type="dataset")
    except Exception as e:  # noqa: BLE001 - no network, bad token, quota ...
        raise RuntimeError(f"could not push {folder} to the Hub as {repo_id}: {type(e).__name__}: {e}") from e
    return f"https://huggingface.co/datasets/{repo_id}"


def main(argv: Optional[List[str]] = None) -> int:
    ap = argparse.ArgumentParser(description="torchvision dataset -> parquet shards + dataset card (+ optional push to the Hugging Face Hub)")
    ap.add_argument("--dataset_name", "--dataset", dest="dataset_name", required=True, choices=SUPPORTED)
    ap.add_argument("--cache_dir", "--root", dest="cache_dir", required=True, help="where torchvision keeps / downloads the raw files")
    ap.add_argument("--out_dir", default=None, help="output folder (default: <cache_dir>/<dataset_name>_parquet)")
    ap.add_argument("--no_download", action="store_true", help="do not try to download: the raw files are already in cache_dir")
    ap.add_argument("--limit", type=int, default=None, help="first N examples of every split (smoke runs)")
    ap.add_argument("--token", default=None, help="Hugging Face token; with --repo_id the folder is pushed to the Hub")
    ap.add_argument("--repo_id", default=None, help="<user or org>/<name> of the dataset repository")
    ap.add_argument("--public", action="store_true")
    ap.add_argument("--dataset_summary", default="")
    ap.add_argument("--dataset_description", default="")
    ap.add_argument("--official_link", default="")
    a = ap.parse_args(argv)
    out = a.out_dir or os.path.join(a.cache_dir, f"{a.dataset_name}_parquet")
    export_parquet(a.dataset_name, a.cache_dir, out, download=not a.no_download, limit=a.limit, summary=a.dataset_summary, description=a.dataset_description, link=a.official_link)
    print(f"wrote {out}/train.parquet, test.parquet, README.md")
    if a.token and a.repo_id:
        try:
            print("pushed to", push_to_hub(out, a.repo_id, a.token, a.public))
        except RuntimeError as e:
            print(e)
            return 2
    elif a.token or a.repo_id:
        print("both --token and --repo_id are needed to push; nothing uploaded")
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
