"""SCAFFOLD client arithmetic on lists of per-layer arrays - shared by the host-side framework bindings (Keras, Flax).

Reference: the Keras callback + optimizer wrapper (p2pfl/learning/frameworks/tensorflow/callbacks/scaffold_callback.py:30-163),
whose formulas these are:  every step ``g <- g + eta_l * (c_i - c)``;  at the end of the local epoch(s)
``c_i <- c_i + (x - y) / (K * eta_l)``, ``delta_y = y - x``, ``delta_c = c_i_new - c_i_old``.
(The device path folds the same terms into the fused optimizer kernel: ``callbacks/scaffold_callback.py``.)
Plain NumPy so the arithmetic is tested on any box, with or without TensorFlow / JAX.
"""

from __future__ import annotations

from typing import Any, List, Optional, Sequence, Tuple

import numpy as np


def corrected_gradients(grads: Sequence[Optional[Any]], c_local: Sequence[Any], c_global: Sequence[Any], eta_l: float) -> List[Optional[Any]]:
    """``g + eta_l * (c_i - c)`` per layer; layers without a gradient (``None``) stay ``None``.  Works on anything with array
    arithmetic (NumPy arrays, tf tensors / variables, jax arrays)."""
    return [None if g is None else g + eta_l * (ci - c) for g, ci, c in zip(grads, c_local, c_global)]


def finalize(x_start: Sequence[np.ndarray], y_end: Sequence[np.ndarray], c_local: Sequence[np.ndarray], steps: int, eta_l: float) -> Tuple[List[np.ndarray], List[np.ndarray], List[np.ndarray]]:
    """After ``steps`` local steps from ``x_start`` to ``y_end``: ``(new c_i, delta_y, delta_c)``."""
    if steps < 1 or eta_l <= 0:
        raise ValueError("SCAFFOLD needs at least one local step and a positive learning rate")
    scale = 1.0 / (steps * eta_l)
    new_c, dy, dc = [], [], []
    for x, y, ci in zip(x_start, y_end, c_local):
        x, y, ci = np.asarray(x), np.asarray(y), np.asarray(ci)
        adj = (x - y) * scale
        new_c.append(ci + adj)
        dy.append(y - x)
        dc.append(adj)
    return new_c, dy, dc
