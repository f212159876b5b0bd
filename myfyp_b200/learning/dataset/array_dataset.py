"""A tiny columnar dataset: dict of equally long NumPy arrays.

The reference uses Hugging Face ``datasets.Dataset`` as its intermediate representation
(p2pfl/learning/dataset/p2pfl_dataset.py:59-99).  HF datasets are still accepted by
:class:`P2PFLDataset`; this class is the zero-dependency, zero-copy backend used for synthetic
and in-memory data (there is no network on the target boxes) and it mimics the handful of
``Dataset`` methods the framework relies on.
"""

from __future__ import annotations

from typing import Any, Dict, Iterable, List, Sequence, Union

import numpy as np


class ArrayDataset:
    def __init__(self, columns: Dict[str, Any]) -> None:
        self._cols = {k: np.asarray(v) for k, v in columns.items()}
        lens = {len(v) for v in self._cols.values()}
        if len(lens) > 1:
            raise ValueError(f"columns have different lengths: {lens}")
        self._len = lens.pop() if lens else 0

    @property
    def column_names(self) -> List[str]:
        return list(self._cols.keys())

    def __len__(self) -> int:
        return self._len

    def __getitem__(self, idx: Union[int, slice, str, Sequence[int]]) -> Any:
        if isinstance(idx, str):
            return self._cols[idx]
        return {k: v[idx] for k, v in self._cols.items()}

    def select(self, indices: Iterable[int]) -> "ArrayDataset":
        ind = np.asarray(list(indices), dtype=np.int64)
        return ArrayDataset({k: v[ind] for k, v in self._cols.items()})

    def train_test_split(self, test_size: float = 0.2, seed: int = 42, shuffle: bool = True, **_: Any) -> "ArrayDatasetDict":
        n = len(self)
        ind = np.arange(n)
        if shuffle:
            np.random.default_rng(seed).shuffle(ind)
        n_test = int(round(n * test_size)) if test_size < 1 else int(test_size)
        n_test = min(max(n_test, 1), n - 1) if n > 1 else 0
        return ArrayDatasetDict({"train": self.select(ind[: n - n_test]), "test": self.select(ind[n - n_test :])})

    def to_dict(self) -> Dict[str, np.ndarray]:
        return dict(self._cols)


class ArrayDatasetDict(dict):
    """split name -> ArrayDataset"""
