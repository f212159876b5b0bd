"""Disk checkpoints in the reference's encoded-parameter layout.

The reference has no disk checkpointing (Lightning's is disabled, lightning_learner.py:86); its only
serialised model form is the wire pickle ``{"params": [ndarray...], "additional_info": {...}}``
(p2pfl_model.py:81-85), which BASELINE.json calls "the encoded-parameter checkpoint layout".  A checkpoint
here is that exact blob (so ``P2PFLModel.decode_parameters`` of either framework reads it) plus a small JSON
side-car with the experiment position, enough to resume.
"""

from __future__ import annotations

import json
import os
from typing import Any, Dict, Optional


def save_checkpoint(path: str, model, round: Optional[int] = None, total_rounds: Optional[int] = None, exp_name: Optional[str] = None, extra: Optional[Dict[str, Any]] = None) -> None:
    os.makedirs(os.path.dirname(os.path.abspath(path)) or ".", exist_ok=True)
    tmp = path + ".tmp"
    with open(tmp, "wb") as f:
        f.write(model.encode_parameters())
    os.replace(tmp, path)
    meta = {
        "round": round, "total_rounds": total_rounds, "exp_name": exp_name,
        "contributors": list(getattr(model, "contributors", [])), "num_samples": int(getattr(model, "num_samples", 0)),
        "format": "p2pfl-encoded-parameters/pickle", "extra": extra or {},
    }
    with open(path + ".json", "w") as f:
        json.dump(meta, f)


def load_checkpoint(path: str, model) -> Dict[str, Any]:
    """Load parameters into ``model`` (raises ``ModelNotMatchingError`` on a different architecture); returns the side-car."""
    with open(path, "rb") as f:
        model.set_parameters(f.read())
    meta: Dict[str, Any] = {}
    if os.path.exists(path + ".json"):
        with open(path + ".json") as f:
            meta = json.load(f)
        if meta.get("contributors") and meta.get("num_samples"):
            model.set_contribution(meta["contributors"], meta["num_samples"])
    return meta


def save_node(node, directory: str) -> str:
    """Checkpoint a node's current model with its experiment position."""
    st = node.state
    path = os.path.join(directory, f"{node.addr.replace(':', '_').replace('/', '_')}.ckpt")
    save_checkpoint(path, node.get_model(), st.round, st.total_rounds, st.exp_name)
    return path


def resume_node(node, path: str) -> int:
    """Load a checkpoint into ``node``'s model and return the number of rounds that were still to run when it was written
    (``node.set_start_learning(rounds=resume_node(node, path), epochs=...)`` on the initiator continues the experiment; the other
    nodes receive the restored parameters through the usual ``init_model`` diffusion)."""
    meta = load_checkpoint(path, node.get_model())
    if node.learner is not None:
        node.learner.set_model(node.get_model())
    total, done = meta.get("total_rounds"), meta.get("round")
    return max(int(total) - int(done), 0) if total is not None and done is not None else 0
