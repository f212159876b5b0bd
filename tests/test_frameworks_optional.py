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


def test_scaffold_client_arithmetic_for_the_host_side_bindings():
    """The formulas of the reference's Keras SCAFFOLD wrapper (tensorflow/callbacks/scaffold_callback.py:49-59, 124-140), which
    the Keras / Flax adaptors here delegate to - checked on NumPy, so no TensorFlow is needed."""
    import numpy as np

    from myfyp_b200.learning.frameworks import scaffold_math as sm

    g = [np.array([1.0, 2.0]), None, np.array([[0.5]])]
    ci = [np.array([0.1, 0.2]), np.array([9.0]), np.array([[0.3]])]
    c = [np.array([0.0, 0.1]), np.array([9.0]), np.array([[0.1]])]
    out = sm.corrected_gradients(g, ci, c, eta_l=0.5)
    np.testing.assert_allclose(out[0], [1.05, 2.05])
    assert out[1] is None
    np.testing.assert_allclose(out[2], [[0.6]])
    x, y = [np.array([1.0, 1.0])], [np.array([0.8, 1.4])]
    new_c, dy, dc = sm.finalize(x, y, [np.array([0.5, 0.5])], steps=4, eta_l=0.1)
    np.testing.assert_allclose(dy[0], [-0.2, 0.4])
    np.testing.assert_allclose(dc[0], [0.5, -1.0])         # (x - y) / (K * eta)
    np.testing.assert_allclose(new_c[0], [1.0, -0.5])      # c_i + adjustment
    import pytest

    with pytest.raises(ValueError):
        sm.finalize(x, y, [np.zeros(2)], steps=0, eta_l=0.1)


def test_flax_wire_form_is_the_reference_s_pickled_dict():
    """flax/flax_model.py:102-127: the blob holds the nested parameter dict; receipt flattens it in insertion order."""
    import pickle

    import numpy as np

    from myfyp_b200.learning.frameworks.flax import wire

    tree = {"Dense_0": {"kernel": np.ones((3, 2), np.float32), "bias": np.zeros(2, np.float32)}, "Dense_1": {"kernel": np.full((2, 1), 2.0, np.float32), "bias": np.array([5.0], np.float32)}}
    blob = wire.encode(tree, {"scaffold": {"k": 1}})
    raw = pickle.loads(blob)
    assert list(raw["params"].keys()) == ["Dense_0", "Dense_1"] and list(raw["params"]["Dense_0"].keys()) == ["kernel", "bias"]
    leaves, info = wire.decode(blob)
    assert [l.shape for l in leaves] == [(3, 2), (2,), (2, 1), (1,)] and info == {"scaffold": {"k": 1}}
    # the reference's own flattening of the same blob gives the same list
    ref_leaves = [np.array(v) for layer in raw["params"].values() for v in layer.values()]
    assert all(np.array_equal(a, b) for a, b in zip(leaves, ref_leaves))
    back = wire.fill_tree(tree, [l + 1 for l in leaves])
    assert float(back["Dense_1"]["bias"][0]) == 6.0
    # the list form of the other bindings is still accepted
    assert len(wire.wire_to_leaves([np.zeros(2), np.ones(3)])) == 2
