"""``FlaxModel``: ``P2PFLModel`` over a ``flax.linen.Module`` + its parameter pytree (reference: flax/flax_model.py:37-163).

The wire format is the same list of arrays as every other binding: the pytree's leaves in ``jax.tree_util`` order (the
reference pickles the dict itself, which only another Flax node can read)."""

from __future__ import annotations

import copy
from typing import Any, Dict, List, Optional, Union

import jax
import numpy as np
from flax import linen as nn

from myfyp_b200.learning.exceptions import ModelNotMatchingError
from myfyp_b200.learning.p2pfl_model import P2PFLModel


class FlaxModel(P2PFLModel):
    def __init__(self, model: nn.Module, init_params: Any = None, params: Optional[Union[List[np.ndarray], bytes]] = None, num_samples: Optional[int] = None,
                 contributors: Optional[List[str]] = None, additional_info: Optional[Dict[str, Any]] = None) -> None:
        if init_params is None:
            raise ValueError("FlaxModel needs the initial parameter pytree (module.init(key, sample)['params']).")
        self.model_params = init_params
        super().__init__(model, params, num_samples, contributors, additional_info)

    def get_parameters(self) -> List[np.ndarray]:
        return [np.asarray(leaf) for leaf in jax.tree_util.tree_leaves(self.model_params)]

    def set_parameters(self, params: Union[List[np.ndarray], bytes]) -> None:
        if isinstance(params, (bytes, bytearray)):
            params, info = self.decode_parameters(bytes(params))
            self.additional_info.update(info)
        leaves, treedef = jax.tree_util.tree_flatten(self.model_params)
        if len(params) != len(leaves) or any(np.shape(a) != np.shape(b) for a, b in zip(params, leaves)):
            raise ModelNotMatchingError("Not matching models")
        self.model_params = jax.tree_util.tree_unflatten(treedef, [jax.numpy.asarray(p, dtype=l.dtype) for p, l in zip(params, leaves)])

    # wire format = the reference's: the pickled parameter dict (flax/flax_model.py:102-127), not the list form of the other bindings
    def encode_parameters(self, params=None) -> bytes:
        from myfyp_b200.learning.frameworks.flax import wire
        from myfyp_b200.learning.p2pfl_model import _info_to_numpy

        tree = self.model_params if params is None else wire.fill_tree(self.model_params, list(params))
        return wire.encode(tree, _info_to_numpy(self.additional_info))

    def decode_parameters(self, data: bytes):
        from myfyp_b200.learning.exceptions import DecodingParamsError
        from myfyp_b200.learning.frameworks.flax import wire

        try:
            return wire.decode(data)
        except Exception as e:  # noqa: BLE001
            raise DecodingParamsError("Error decoding parameters") from e

    def build_copy(self, **kwargs) -> "FlaxModel":
        return FlaxModel(self.model, init_params=copy.deepcopy(self.model_params), **kwargs)

    def get_framework(self) -> str:
        return "flax"


class MLP(nn.Module):
    """Dense(256, relu) - Dense(128, relu) - Dense(10) (logits) on flattened inputs."""

    hidden: tuple = (256, 128)
    out_channels: int = 10

    @nn.compact
    def __call__(self, x):
        x = x.reshape((x.shape[0], -1))
        for h in self.hidden:
            x = nn.relu(nn.Dense(h)(x))
        return nn.Dense(self.out_channels)(x)
