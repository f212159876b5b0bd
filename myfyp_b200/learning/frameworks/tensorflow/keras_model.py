"""``KerasModel``: ``P2PFLModel`` over a *built* ``tf.keras.Model`` (reference: tensorflow/keras_model.py:44-113) and the
example ``MLP`` (784-256-128-10, same as the PyTorch one)."""

from __future__ import annotations

import copy
from typing import Any, Dict, List, Optional, Union

import numpy as np
import tensorflow as tf  # noqa: F401 - ImportError here is the signal that the binding is unavailable

from myfyp_b200.learning.exceptions import ModelNotMatchingError
from myfyp_b200.learning.p2pfl_model import P2PFLModel


class KerasModel(P2PFLModel):
    def __init__(self, model: "tf.keras.Model", params: Optional[Union[List[np.ndarray], bytes]] = None, num_samples: Optional[int] = None,
                 contributors: Optional[List[str]] = None, additional_info: Optional[Dict[str, Any]] = None) -> None:
        if not model.built or len(model.get_weights()) == 0:
            raise ValueError("The Keras model must be built (call it once or pass input_shape) before it is wrapped.")
        super().__init__(model, params, num_samples, contributors, additional_info)

    def get_parameters(self) -> List[np.ndarray]:
        return [np.asarray(w) for w in self.model.get_weights()]

    def set_parameters(self, params: Union[List[np.ndarray], bytes]) -> None:
        if isinstance(params, (bytes, bytearray)):
            params, info = self.decode_parameters(bytes(params))
            self.additional_info.update(info)
        current = self.model.get_weights()
        if len(params) != len(current) or any(np.shape(a) != np.shape(b) for a, b in zip(params, current)):
            raise ModelNotMatchingError("Not matching models")
        self.model.set_weights([np.asarray(p) for p in params])

    def build_copy(self, **kwargs) -> "KerasModel":
        clone = tf.keras.models.clone_model(self.model)
        clone.build(self.model.input_shape)
        clone.set_weights(copy.deepcopy(self.model.get_weights()))
        return KerasModel(clone, **kwargs)

    def get_framework(self) -> str:
        return "tensorflow"


def MLP(input_shape=(28, 28), hidden=(256, 128), out_channels: int = 10) -> "tf.keras.Model":
    """Flatten - Dense(256, relu) - Dense(128, relu) - Dense(10) (logits), built."""
    layers = [tf.keras.layers.Input(shape=tuple(input_shape)), tf.keras.layers.Flatten()]
    layers += [tf.keras.layers.Dense(h, activation="relu") for h in hidden]
    layers.append(tf.keras.layers.Dense(out_channels))
    return tf.keras.Sequential(layers)
