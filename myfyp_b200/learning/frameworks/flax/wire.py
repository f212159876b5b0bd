"""Wire form of a Flax parameter tree.  The reference's ``FlaxModel`` does not ship the common list of arrays but pickles the
parameter DICT itself: ``{"params": {layer: {name: array}}, "additional_info": {...}}`` (p2pfl/learning/frameworks/flax/
flax_model.py:102-127) and flattens it on receipt in insertion order (``__dict_to_np``, :60-61).  These helpers produce and read
that form (and still accept the list form every other binding uses); no JAX needed, so they are tested everywhere."""

from __future__ import annotations

import pickle
from typing import Any, Dict, List, Tuple

import numpy as np


def tree_to_wire(tree: Any) -> Any:
    """Nested mapping of arrays -> the same nesting with plain NumPy leaves (insertion order kept)."""
    if isinstance(tree, dict) or hasattr(tree, "items"):
        return {k: tree_to_wire(v) for k, v in tree.items()}
    return np.asarray(tree)


def wire_to_leaves(obj: Any) -> List[np.ndarray]:
    """Leaves in the reference's flattening order (values of the outer dict, then of the inner ones); a list passes through."""
    if isinstance(obj, dict) or hasattr(obj, "items"):
        out: List[np.ndarray] = []
        for v in obj.values():
            out.extend(wire_to_leaves(v))
        return out
    if isinstance(obj, (list, tuple)):
        return [np.asarray(a) for a in obj]
    return [np.asarray(obj)]


def fill_tree(template: Any, leaves: List[np.ndarray]) -> Any:
    """A copy of ``template``'s nesting holding ``leaves`` (same order as :func:`wire_to_leaves`)."""
    it = iter(leaves)

    def go(node):
        if isinstance(node, dict) or hasattr(node, "items"):
            return {k: go(v) for k, v in node.items()}
        return next(it)

    out = go(template)
    if next(it, None) is not None:
        raise ValueError("more parameters than the model has")
    return out


def encode(tree: Any, additional_info: Dict[str, Any]) -> bytes:
    return pickle.dumps({"params": tree_to_wire(tree), "additional_info": additional_info})


def decode(data: bytes) -> Tuple[List[np.ndarray], Dict[str, Any]]:
    loaded = pickle.loads(data)
    return wire_to_leaves(loaded["params"]), loaded.get("additional_info", {})
