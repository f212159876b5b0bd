"""The framework seam (reference: learning/frameworks/__init__.py:14-19 and learner_factory.py:29-56).

The reference dispatches on ``model.get_framework()`` to a Lightning, a Keras or a Flax learner.  This framework is
PyTorch-on-B200: PyTorch is the layer the hand-written kernels plug into.  The Keras and Flax bindings are thin host-side
wrappers (``tensorflow/``, ``flax/``; importable only where those frameworks are installed) so that models a user already has
there can join a federation through the wire format; a user's own ``Learner`` subclass can be registered for any name.
The sub-packages mirror the reference's module paths (``frameworks.pytorch.lightning_model`` ...).
"""

from __future__ import annotations

import enum
from typing import Callable, Dict


class Framework(enum.Enum):
    PYTORCH = "pytorch"
    TENSORFLOW = "tensorflow"
    FLAX = "flax"


_REGISTRY: Dict[str, Callable[[], type]] = {}


def register_learner(framework: str, loader: Callable[[], type]) -> None:
    """``loader()`` returns the ``Learner`` subclass for models whose ``get_framework()`` is ``framework``."""
    _REGISTRY[str(framework)] = loader


def learner_for(framework: str) -> type:
    if framework in _REGISTRY:
        return _REGISTRY[framework]()
    known = ", ".join(sorted(_REGISTRY)) or "none"
    raise ValueError(
        f"No learner is registered for framework {framework!r} (registered: {known}); "
        "register one with myfyp_b200.learning.frameworks.register_learner()."
    )


def _torch_learner() -> type:
    from myfyp_b200.learning.torch_learner import TorchLearner

    return TorchLearner


def _keras_learner() -> type:
    from myfyp_b200.learning.frameworks.tensorflow.keras_learner import KerasLearner

    return KerasLearner


def _flax_learner() -> type:
    from myfyp_b200.learning.frameworks.flax.flax_learner import FlaxLearner

    return FlaxLearner


register_learner(Framework.PYTORCH.value, _torch_learner)
# host-side bindings for models a user already has in another framework (they need tensorflow resp. jax / flax / optax to be
# installed; the B200 kernels are reached through the PyTorch binding only)
register_learner(Framework.TENSORFLOW.value, _keras_learner)
register_learner(Framework.FLAX.value, _flax_learner)
