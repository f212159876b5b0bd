"""``KerasExportStrategy``: a split as a batched ``tf.data.Dataset`` (reference: tensorflow/keras_dataset.py:29-69, batch 1 there)."""

from __future__ import annotations

from typing import Any, Callable, Optional

import numpy as np
import tensorflow as tf

from myfyp_b200.learning.dataset.p2pfl_dataset import DataExportStrategy


class KerasExportStrategy(DataExportStrategy):
    @staticmethod
    def export(data: Any, transforms: Optional[Callable] = None, batch_size: int = 32, image_key: str = "image", label_key: str = "label", **kwargs):
        if hasattr(data, "with_format"):  # HF Dataset
            d = data.with_format("numpy")
            x, y = np.asarray(d[image_key]), np.asarray(d[label_key])
        else:
            x, y = np.asarray(data[image_key]), np.asarray(data[label_key])
        if transforms is not None:
            out = transforms({image_key: x, label_key: y})
            x, y = out[image_key], out[label_key]
        return tf.data.Dataset.from_tensor_slices((x.astype("float32"), y.astype("int64"))).batch(batch_size)
