"""The framework seam: reference module paths resolve; the Keras / Flax bindings run where those frameworks exist
(they are not part of the build image, so those tests skip there)."""

import importlib

import numpy as np
import pytest

from myfyp_b200.learning.dataset import P2PFLDataset


def test_reference_module_paths_resolve():
    from myfyp_b200.learning.frameworks import Framework, learner_for
    from myfyp_b200.learning.frameworks.callback_factory import CallbackFactory  # noqa: F401
    from myfyp_b200.learning.frameworks.exceptions import DecodingParamsError, ModelNotMatchingError  # noqa: F401
    from myfyp_b200.learning.frameworks.learner_factory import LearnerFactory
    from myfyp_b200.learning.frameworks.p2pfl_model import P2PFLModel  # noqa: F401
    from myfyp_b200.learning.frameworks.pytorch.lightning_dataset import PyTorchExportStrategy  # noqa: F401
    from myfyp_b200.learning.frameworks.pytorch.lightning_learner import LightningLearner
    from myfyp_b200.learning.frameworks.pytorch.lightning_model import MLP, LightningModel
    from myfyp_b200.models import mlp_model

    assert {f.value for f in Framework} == {"pytorch", "tensorflow", "flax"}
    assert LearnerFactory.create_learner(mlp_model(seed=0)) is LightningLearner is learner_for("pytorch")
    assert LightningModel(MLP()).get_framework() == "pytorch"
    for fw, dep in (("tensorflow", "tensorflow"), ("flax", "jax")):
        try:
            importlib.import_module(dep)
        except ImportError:
            with pytest.raises(ImportError):  # registered, but its framework is not installed: a clear import error
                learner_for(fw)


def _tiny():
    return P2PFLDataset.synthetic_mnist(256, 64, seed=3)


def test_keras_binding_round_trip_and_fit():
    pytest.importorskip("tensorflow")
    from myfyp_b200.learning.exceptions import ModelNotMatchingError
    from myfyp_b200.learning.frameworks.tensorflow.keras_learner import KerasLearner
    from myfyp_b200.learning.frameworks.tensorflow.keras_model import MLP, KerasModel

    m = KerasModel(MLP())
    blob = m.encode_parameters()
    other = m.build_copy(params=blob, num_samples=3, contributors=["x"])
    for a, b in zip(m.get_parameters(), other.get_parameters()):
        np.testing.assert_array_equal(a, b)
    with pytest.raises(ModelNotMatchingError):
        m.set_parameters([np.zeros((2, 2))])
    learner = KerasLearner(m, _tiny(), "k0", batch_size=32)
    learner.set_epochs(3)
    before = learner.evaluate()["test_loss"]
    learner.fit()
    assert learner.evaluate()["test_loss"] < before and m.get_contributors() == ["k0"]


def test_flax_binding_round_trip_and_fit():
    pytest.importorskip("jax")
    pytest.importorskip("flax")
    pytest.importorskip("optax")
    import jax

    from myfyp_b200.learning.exceptions import ModelNotMatchingError
    from myfyp_b200.learning.frameworks.flax.flax_learner import FlaxLearner
    from myfyp_b200.learning.frameworks.flax.flax_model import MLP, FlaxModel

    module = MLP()
    params = module.init(jax.random.PRNGKey(0), np.zeros((1, 28, 28), np.float32))["params"]
    m = FlaxModel(module, init_params=params)
    other = m.build_copy(params=m.encode_parameters(), num_samples=3, contributors=["x"])
    for a, b in zip(m.get_parameters(), other.get_parameters()):
        np.testing.assert_array_equal(a, b)
    with pytest.raises(ModelNotMatchingError):
        m.set_parameters([np.zeros((2, 2))])
    learner = FlaxLearner(m, _tiny(), "f0", batch_size=32)
    learner.set_epochs(3)
    before = learner.evaluate()["test_loss"]
    learner.fit()
    assert learner.evaluate()["test_loss"] < before and m.get_contributors() == ["f0"]
