"""``FlaxExportStrategy``: a split as a list of ``(x, y)`` NumPy batches (reference: flax/flax_dataset.py:30-67)."""

from __future__ import annotations

from typing import Any, Callable, List, Optional, Tuple

import numpy as np

from myfyp_b200.learning.dataset.p2pfl_dataset import DataExportStrategy


class FlaxExportStrategy(DataExportStrategy):
    @staticmethod
    def export(data: Any, transforms: Optional[Callable] = None, batch_size: int = 32, image_key: str = "image", label_key: str = "label", **kwargs) -> List[Tuple[np.ndarray, np.ndarray]]:
        if hasattr(data, "with_format"):  # HF Dataset
            d = data.with_format("numpy")
            x, y = np.asarray(d[image_key]), np.asarray(d[label_key])
        else:
            x, y = np.asarray(data[image_key]), np.asarray(data[label_key])
        if transforms is not None:
            out = transforms({image_key: x, label_key: y})
            x, y = out[image_key], out[label_key]
        x, y = x.astype("float32"), y.astype("int32")
        return [(x[i : i + batch_size], y[i : i + batch_size]) for i in range(0, len(x), batch_size)]
