"""How to split one dataset among N federated nodes.

Reference: p2pfl/learning/dataset/partition_strategies.py:33-436.  Same interface -
``Strategy.generate_partitions(train, test, num_partitions, seed=666, label_tag="label", **kw)``
returns ``(train index lists, test index lists)``.  ``RandomIID`` and ``Dirichlet`` follow the
reference's behaviour; ``LabelSkewed`` (reference raises ``NotImplementedError`` at :142) and
``PercentageBasedNonIID`` (reference: empty class, :433-436) are implemented here.
"""

from __future__ import annotations

import abc
import random
from typing import Any, List, Tuple

import numpy as np


def _labels(data: Any, label_tag: str) -> np.ndarray:
    return np.asarray(data[label_tag])


class DataPartitionStrategy(abc.ABC):
    @staticmethod
    @abc.abstractmethod
    def generate_partitions(train_data: Any, test_data: Any, num_partitions: int, **kwargs) -> Tuple[List[List[int]], List[List[int]]]: ...


class RandomIIDPartitionStrategy(DataPartitionStrategy):
    """Seeded shuffle, equal split, the first ``len % n`` partitions get one extra sample."""

    @staticmethod
    def generate_partitions(train_data, test_data, num_partitions: int, seed: int = 666, **kwargs):
        return (_iid(len(train_data), seed, num_partitions), _iid(len(test_data), seed, num_partitions))


def _iid(n: int, seed: int, parts: int) -> List[List[int]]:
    idx = list(range(n))
    random.Random(seed).shuffle(idx)
    base, rem = divmod(n, parts)
    return [idx[i * base + min(i, rem) : (i + 1) * base + min(i + 1, rem)] for i in range(parts)]


class LabelSkewedPartitionStrategy(DataPartitionStrategy):
    """Sort by label and deal contiguous shards: each node sees few classes (pathological non-IID)."""

    @staticmethod
    def generate_partitions(train_data, test_data, num_partitions: int, seed: int = 666, label_tag: str = "label", shards_per_node: int = 2, **kwargs):
        def split(data):
            y = _labels(data, label_tag)
            order = np.argsort(y, kind="stable")
            shards = np.array_split(order, num_partitions * shards_per_node)
            perm = np.random.default_rng(seed).permutation(len(shards))
            return [sorted(np.concatenate([shards[perm[i * shards_per_node + j]] for j in range(shards_per_node)]).tolist()) for i in range(num_partitions)]

        return split(train_data), split(test_data)


class PercentageBasedNonIIDPartitionStrategy(DataPartitionStrategy):
    """``iid_fraction`` of every node's share is drawn IID, the rest comes from label-sorted shards."""

    @staticmethod
    def generate_partitions(train_data, test_data, num_partitions: int, seed: int = 666, label_tag: str = "label", iid_fraction: float = 0.5, **kwargs):
        def split(data):
            n = len(data)
            rng = np.random.default_rng(seed)
            perm = rng.permutation(n)
            n_iid = int(n * iid_fraction)
            iid_part, rest = perm[:n_iid], perm[n_iid:]
            y = _labels(data, label_tag)
            rest = rest[np.argsort(y[rest], kind="stable")]
            a, b = np.array_split(iid_part, num_partitions), np.array_split(rest, num_partitions)
            return [sorted(np.concatenate([a[i], b[i]]).tolist()) for i in range(num_partitions)]

        return split(train_data), split(test_data)


class DirichletPartitionStrategy(DataPartitionStrategy):
    """Label-distribution skew: class c is spread over the nodes with proportions ~ Dir(alpha).

    ``alpha``: float or one value per partition (small = more skew); ``min_partition_size``: resample (at
    most ``max_tries`` times) until every partition has at least that many samples;
    ``self_balancing``: a partition that already holds its fair share stops receiving samples.
    """

    @staticmethod
    def generate_partitions(
        train_data,
        test_data,
        num_partitions: int,
        seed: int = 666,
        label_tag: str = "label",
        alpha: Any = 0.1,
        min_partition_size: int = 2,
        self_balancing: bool = False,
        max_tries: int = 10,
        **kwargs,
    ):
        a = DirichletPartitionStrategy._alpha(alpha, num_partitions)

        def split(data, min_size):
            y = _labels(data, label_tag)
            n = len(y)
            classes = np.unique(y)
            rng = np.random.default_rng(seed)
            fair = n / num_partitions
            parts: List[List[int]] = []
            for _ in range(max_tries):
                parts = [[] for _ in range(num_partitions)]
                for c in classes:
                    idx = np.flatnonzero(y == c)
                    rng.shuffle(idx)
                    p = rng.dirichlet(a)
                    if self_balancing:
                        p = np.array([pi if len(parts[i]) < fair else 0.0 for i, pi in enumerate(p)])
                        p = p / p.sum() if p.sum() > 0 else np.full(num_partitions, 1.0 / num_partitions)
                    cuts = (np.cumsum(p) * len(idx)).astype(int)[:-1]
                    for i, chunk in enumerate(np.split(idx, cuts)):
                        parts[i].extend(chunk.tolist())
                if min(len(p_) for p_ in parts) >= min_size:
                    break
            else:
                raise ValueError(f"could not reach min_partition_size={min_size} in {max_tries} tries; raise alpha or lower the minimum")
            for p_ in parts:
                rng.shuffle(p_)
            return parts

        return split(train_data, min_partition_size), split(test_data, min(min_partition_size, max(len(test_data) // max(num_partitions, 1) // 4, 0)))

    @staticmethod
    def _alpha(alpha: Any, n: int) -> np.ndarray:
        if isinstance(alpha, (int, float)):
            a = np.full(n, float(alpha))
        else:
            a = np.asarray(alpha, dtype=np.float64)
            if a.shape != (n,):
                raise ValueError("alpha must be a number or have one entry per partition")
        if (a <= 0).any():
            raise ValueError("alpha must be strictly positive")
        return a
