"""Network topologies (reference: p2pfl/utils/topologies.py:31-93).

``STAR / FULL / LINE / RING`` as in the reference plus ``RANDOM`` (seeded Erdos-Renyi made connected by a
random spanning tree) which the reference only lists as planned (docs/source/tutorials/topologies.md:26).
"""

from __future__ import annotations

import time
from enum import Enum
from typing import List, Optional, Union

import numpy as np


class TopologyType(Enum):
    STAR = "star"
    FULL = "full"
    LINE = "line"
    RING = "ring"
    RANDOM = "random"


class TopologyFactory:
    @staticmethod
    def generate_matrix(topology_type: Union[TopologyType, str], num_nodes: int, p: float = 0.3, seed: Optional[int] = None) -> np.ndarray:
        if isinstance(topology_type, str):
            try:
                topology_type = TopologyType(topology_type)
            except ValueError as e:
                raise ValueError(f"Unsupported topology type: {topology_type}") from e
        if not isinstance(topology_type, TopologyType):
            raise TypeError("topology_type must be a TopologyType")
        m = np.zeros((num_nodes, num_nodes), dtype=int)
        if num_nodes < 2:
            return m
        if topology_type == TopologyType.STAR:
            m[0, 1:] = 1
            m[1:, 0] = 1
        elif topology_type == TopologyType.FULL:
            m[:] = 1
            np.fill_diagonal(m, 0)
        elif topology_type == TopologyType.LINE:
            idx = np.arange(num_nodes - 1)
            m[idx, idx + 1] = 1
            m[idx + 1, idx] = 1
        elif topology_type == TopologyType.RING:
            idx = np.arange(num_nodes)
            m[idx, (idx + 1) % num_nodes] = 1
            m[(idx + 1) % num_nodes, idx] = 1
        elif topology_type == TopologyType.RANDOM:
            rng = np.random.default_rng(seed)
            order = rng.permutation(num_nodes)
            for i in range(1, num_nodes):  # random spanning tree => connected
                j = order[rng.integers(0, i)]
                m[order[i], j] = m[j, order[i]] = 1
            extra = np.triu(rng.random((num_nodes, num_nodes)) < p, 1)
            m |= (extra | extra.T).astype(int)
            np.fill_diagonal(m, 0)
        return m

    @staticmethod
    def connect_nodes(adjacency_matrix: np.ndarray, nodes: List, pause: float = 0.0) -> None:
        """Connect ``nodes[i] -> nodes[j]`` for every edge with ``i < j`` (the handshake makes it symmetric)."""
        n = len(nodes)
        for i in range(n):
            for j in range(i + 1, n):
                if adjacency_matrix[i, j] == 1:
                    nodes[i].connect(nodes[j].addr)
                    if pause > 0:
                        time.sleep(pause)  # the reference sleeps 0.1 s per edge
