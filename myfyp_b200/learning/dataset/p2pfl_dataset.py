"""``P2PFLDataset``: split-aware dataset wrapper + export strategies.

Reference: p2pfl/learning/dataset/p2pfl_dataset.py:55-342 (same constructor, ``get``,
``set_transforms``, ``generate_train_test_split``, ``get_num_samples``, ``generate_partitions``,
``export``, ``from_csv/json/parquet/pandas/huggingface/generator``).  The backing store is either a
Hugging Face ``Dataset``/``DatasetDict`` (as in the reference) or the zero-dependency
:class:`ArrayDataset`.  Fixed: ``generate_train_test_split`` honours its arguments.

B200 additions: ``from_arrays`` / ``synthetic_mnist`` / ``synthetic_cifar10`` (no network on the
boxes) and ``TensorExportStrategy`` which hands the learner whole-split tensors so that the
training set lives in HBM and batches are index slices (no DataLoader, no per-batch H2D).
"""

from __future__ import annotations

import abc
from typing import Any, Callable, Dict, Iterable, List, Optional, Type, Union

import numpy as np

from myfyp_b200.learning.dataset.array_dataset import ArrayDataset, ArrayDatasetDict
from myfyp_b200.learning.dataset.partition_strategies import DataPartitionStrategy


class DataExportStrategy(abc.ABC):
    @staticmethod
    @abc.abstractmethod
    def export(data: Any, transforms: Optional[Callable] = None, **kwargs) -> Any: ...


def _is_dict_like(data: Any) -> bool:
    return isinstance(data, dict) or type(data).__name__ == "DatasetDict"


class P2PFLDataset:
    def __init__(self, data: Any, train_split_name: str = "train", test_split_name: str = "test", transforms: Optional[Callable] = None) -> None:
        if isinstance(data, dict) and not isinstance(data, ArrayDatasetDict) and type(data).__name__ != "DatasetDict":
            # {"train": {...columns...}, "test": {...}} or plain columns
            if data and all(isinstance(v, dict) for v in data.values()):
                data = ArrayDatasetDict({k: ArrayDataset(v) for k, v in data.items()})
            else:
                data = ArrayDataset(data)
        self._data = data
        self._train_split_name = train_split_name
        self._test_split_name = test_split_name
        self._transforms = transforms

    # ------------------------------------------------------------------ access
    def _split(self, train: bool) -> Any:
        if not _is_dict_like(self._data):
            raise ValueError("Unsplit dataset: call generate_train_test_split() first")
        return self._data[self._train_split_name if train else self._test_split_name]

    def get(self, idx, train: bool = True) -> Dict[str, Any]:
        item = self._split(train)[idx]
        return self._transforms(item) if self._transforms is not None else item

    def set_transforms(self, transforms: Callable) -> None:
        self._transforms = transforms

    def generate_train_test_split(self, test_size: float = 0.2, seed: int = 42, shuffle: bool = True, **kwargs) -> None:
        if _is_dict_like(self._data):
            raise ValueError("Dataset is already split")
        self._data = self._data.train_test_split(test_size=test_size, seed=seed, shuffle=shuffle, **kwargs)
        self._train_split_name, self._test_split_name = "train", "test"

    def get_num_samples(self, train: bool = True) -> int:
        return len(self._split(train))

    def generate_partitions(self, num_partitions: int, strategy: Union[DataPartitionStrategy, Type[DataPartitionStrategy]], seed: int = 666, label_tag: str = "label", **kwargs) -> List["P2PFLDataset"]:
        train, test = self._split(True), self._split(False)
        tr_idx, te_idx = strategy.generate_partitions(train, test, num_partitions, seed=seed, label_tag=label_tag, **kwargs)
        cls_dict = ArrayDatasetDict if isinstance(train, ArrayDataset) else type(self._data)
        return [
            P2PFLDataset(
                cls_dict({self._train_split_name: train.select(tr_idx[i]), self._test_split_name: test.select(te_idx[i])}),
                self._train_split_name,
                self._test_split_name,
                self._transforms,
            )
            for i in range(num_partitions)
        ]

    def export(self, strategy: Union[DataExportStrategy, Type[DataExportStrategy]], train: bool = True, **kwargs) -> Any:
        return strategy.export(self._split(train), transforms=self._transforms, **kwargs)

    # ------------------------------------------------------------------ constructors (HF-backed, like the reference)
    @classmethod
    def from_csv(cls, data_files, **kwargs) -> "P2PFLDataset":
        from datasets import load_dataset

        return cls(load_dataset("csv", data_files=data_files, **kwargs))

    @classmethod
    def from_json(cls, data_files, **kwargs) -> "P2PFLDataset":
        from datasets import load_dataset

        return cls(load_dataset("json", data_files=data_files, **kwargs))

    @classmethod
    def from_parquet(cls, data_files, **kwargs) -> "P2PFLDataset":
        from datasets import load_dataset

        return cls(load_dataset("parquet", data_files=data_files, **kwargs))

    @classmethod
    def from_pandas(cls, df) -> "P2PFLDataset":
        return cls(ArrayDataset({c: df[c].to_numpy() for c in df.columns}))

    @classmethod
    def from_huggingface(cls, dataset_name: str, **kwargs) -> "P2PFLDataset":
        from datasets import load_dataset

        return cls(load_dataset(dataset_name, **kwargs))

    @classmethod
    def from_generator(cls, generator: Callable[[], Iterable[Dict[str, Any]]]) -> "P2PFLDataset":
        rows = list(generator())
        cols = {k: np.asarray([r[k] for r in rows]) for k in (rows[0].keys() if rows else [])}
        return cls(ArrayDataset(cols))

    # ------------------------------------------------------------------ in-memory / synthetic
    @classmethod
    def from_arrays(cls, train: Dict[str, Any], test: Optional[Dict[str, Any]] = None) -> "P2PFLDataset":
        if test is None:
            return cls(ArrayDataset(train))
        return cls(ArrayDatasetDict({"train": ArrayDataset(train), "test": ArrayDataset(test)}))

    @classmethod
    def synthetic(cls, shape, num_classes: int, n_train: int, n_test: int, seed: int = 0, separable: bool = True) -> "P2PFLDataset":
        """Class-conditional Gaussian blobs of the given sample shape (learnable, so accuracy tests mean something)."""
        rng = np.random.default_rng(seed)
        d = int(np.prod(shape))
        protos = rng.standard_normal((num_classes, d)).astype(np.float32)

        def make(n):
            y = rng.integers(0, num_classes, size=n)
            x = rng.standard_normal((n, d)).astype(np.float32) * 0.8
            if separable:
                x += protos[y]
            return {"image": x.reshape((n, *shape)), "label": y.astype(np.int64)}

        return cls.from_arrays(make(n_train), make(n_test))

    @classmethod
    def synthetic_mnist(cls, n_train: int = 6000, n_test: int = 1000, seed: int = 0) -> "P2PFLDataset":
        return cls.synthetic((28, 28), 10, n_train, n_test, seed)

    @classmethod
    def synthetic_cifar10(cls, n_train: int = 5000, n_test: int = 1000, seed: int = 0) -> "P2PFLDataset":
        return cls.synthetic((3, 32, 32), 10, n_train, n_test, seed)


class TensorExportStrategy(DataExportStrategy):
    """Whole split as ``(x: float32 tensor, y: int64 tensor)`` on ``device`` (pinned host staging when CUDA)."""

    @staticmethod
    def export(data: Any, transforms: Optional[Callable] = None, device: Any = "cpu", image_key: str = "image", label_key: str = "label", **kwargs):
        import torch

        if isinstance(data, ArrayDataset):
            x, y = data[image_key], data[label_key]
        else:  # HF Dataset
            d = data.with_format("numpy")
            x, y = np.asarray(d[image_key]), np.asarray(d[label_key])
        if transforms is not None:
            out = transforms({image_key: x, label_key: y})
            x, y = out[image_key], out[label_key]
        xt = torch.from_numpy(np.ascontiguousarray(x)).float()
        yt = torch.from_numpy(np.ascontiguousarray(y)).long()
        dev = torch.device(device)
        if dev.type == "cuda":
            xt, yt = xt.pin_memory().to(dev, non_blocking=True), yt.pin_memory().to(dev, non_blocking=True)
        return xt, yt


class PyTorchExportStrategy(DataExportStrategy):
    """``torch.utils.data.DataLoader`` over the split (reference: pytorch/lightning_dataset.py:77-103).

    The reference's default ``batch_size`` is 1; it is explicit here and defaults to 32.
    """

    @staticmethod
    def export(data: Any, transforms: Optional[Callable] = None, batch_size: int = 32, shuffle: bool = False, num_workers: int = 0, **kwargs):
        import torch
        from torch.utils.data import DataLoader, Dataset

        class _Wrapped(Dataset):
            def __len__(self_inner) -> int:
                return len(data)

            def __getitem__(self_inner, i):
                item = data[int(i)]
                item = transforms(item) if transforms is not None else item
                return {k: torch.as_tensor(np.asarray(v)) for k, v in item.items()}

        return DataLoader(_Wrapped(), batch_size=batch_size, shuffle=shuffle, num_workers=num_workers)
