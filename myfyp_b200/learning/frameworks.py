"""The framework seam (reference: learning/frameworks/__init__.py:14-19 and learner_factory.py:29-56).

The reference dispatches on ``model.get_framework()`` to a Lightning, a Keras or a Flax learner.  This framework is
PyTorch-on-B200: PyTorch is the layer the hand-written kernels plug into, so only that binding is built.  The enum and
the dispatch are kept so that a model wrapper announcing another framework fails with a clear message instead of an
``ImportError`` somewhere deep, and so that a user's own ``Learner`` subclass can be registered for any name.
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
        f"No learner is registered for framework {framework!r} (registered: {known}).  The TensorFlow / Keras and Flax bindings of "
        "the reference are not built here; register your own with myfyp_b200.learning.frameworks.register_learner()."
    )


def _torch_learner() -> type:
    from myfyp_b200.learning.torch_learner import TorchLearner

    return TorchLearner


register_learner(Framework.PYTORCH.value, _torch_learner)
