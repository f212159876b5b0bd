"""Model wrapper, wire format, datasets, topologies, checkpoints (reference: test/learning/frameworks_test.py,
p2pfl_dataset_test.py, test/utils_test.py)."""

import os
import pickle
import sys

import numpy as np
import pytest
import torch

from myfyp_b200.learning.dataset import (
    DirichletPartitionStrategy,
    LabelSkewedPartitionStrategy,
    P2PFLDataset,
    PercentageBasedNonIIDPartitionStrategy,
    PyTorchExportStrategy,
    RandomIIDPartitionStrategy,
    TensorExportStrategy,
)
from myfyp_b200.learning.exceptions import DecodingParamsError, ModelNotMatchingError
from myfyp_b200.models import cnn_model, mlp_model, resnet18_model
from myfyp_b200.utils.checkpoint import load_checkpoint, save_checkpoint
from myfyp_b200.utils.topologies import TopologyFactory, TopologyType


def test_param_roundtrip_and_slab_backing():
    m = mlp_model(seed=0)
    params = m.get_parameters()
    assert [p.shape for p in params] == [(256, 784), (256,), (128, 256), (128,), (10, 128), (10,)]
    assert sum(p.size for p in params) == 235146
    m.set_parameters([p + 1 for p in params])
    for a, b in zip(m.get_parameters(), params):
        np.testing.assert_allclose(a, b + 1)
    # the nn.Module really lives in the slab: a write through the module is visible in the flat buffer
    w = m.get_model().layers[0].weight
    with torch.no_grad():
        w.fill_(7.0)
    e = m.layout.entries["layers.0.weight"]
    assert torch.all(m.slab[e.offset : e.offset + e.numel] == 7.0)
    assert w.grad is not None and w.grad.data_ptr() == m.grad_slab[e.offset : e.offset + 1].data_ptr()


def test_encode_decode_is_the_reference_wire_format():
    m = mlp_model(seed=1)
    m.add_info("cb", {"k": 3})
    blob = m.encode_parameters()
    raw = pickle.loads(blob)
    assert set(raw.keys()) == {"params", "additional_info"} and raw["additional_info"] == {"cb": {"k": 3}}
    assert all(isinstance(p, np.ndarray) and p.dtype == np.float32 for p in raw["params"])
    m2 = mlp_model(seed=2)
    m2.set_parameters(blob)
    for a, b in zip(m.get_parameters(), m2.get_parameters()):
        np.testing.assert_array_equal(a, b)
    assert m2.get_info("cb") == {"k": 3}
    assert m2.encode_parameters() == blob  # byte-identical re-encode
    with pytest.raises(DecodingParamsError):
        m2.set_parameters(b"not a pickle")


def test_reference_can_decode_our_blob():
    """Interop: the reference's own P2PFLModel.decode_parameters reads what we emit (and we read what it emits)."""
    ref_root = "/root/reference"
    if not os.path.isdir(os.path.join(ref_root, "p2pfl")):
        pytest.skip("reference tree not mounted")
    sys.path.insert(0, ref_root)
    try:
        cwd = os.getcwd()
        import tempfile

        os.chdir(tempfile.mkdtemp())  # the reference's logger writes ./logs
        from p2pfl.learning.frameworks.p2pfl_model import P2PFLModel as RefModel
    except Exception as e:  # noqa: BLE001
        pytest.skip(f"reference not importable: {e}")
    finally:
        os.chdir(cwd)
        sys.path.remove(ref_root)
    m = mlp_model(seed=3)
    params, info = RefModel(None).decode_parameters(m.encode_parameters())
    for a, b in zip(params, m.get_parameters()):
        np.testing.assert_array_equal(a, b)
    blob = RefModel(None).encode_parameters(params=[p * 2 for p in params])
    m.set_parameters(blob)
    np.testing.assert_allclose(m.get_parameters()[0], params[0] * 2)


def test_wrong_architecture_raises():
    m = mlp_model(seed=0)
    other = cnn_model(seed=0)
    with pytest.raises(ModelNotMatchingError):
        m.set_parameters(other.get_parameters())
    with pytest.raises(ModelNotMatchingError):
        m.set_parameters(other.encode_parameters())
    with pytest.raises(ModelNotMatchingError):
        m.build_copy(params=other.publish(), num_samples=1, contributors=["x"])


def test_build_copy_is_held_and_zero_copy_for_descriptors():
    m = mlp_model(seed=0)
    ref = m.publish()
    held = m.build_copy(params=ref, num_samples=5, contributors=["a"])
    assert held.get_model() is None and held.slab.data_ptr() == ref.tensor.data_ptr()
    with torch.no_grad():
        m.slab.add_(1.0)  # the owner keeps training: the published snapshot must not move
    np.testing.assert_allclose(held.get_parameters()[0] + 1, m.get_parameters()[0], atol=1e-6)
    assert held.get_contributors() == ["a"] and held.get_num_samples() == 5
    with pytest.raises(ValueError):
        m.build_copy().get_contributors()


def test_resnet_int_buffers_average_then_cast():
    m = resnet18_model(seed=0, width=8)
    params = m.get_parameters()
    names = m.layout.order
    i = names.index("bn1.num_batches_tracked")
    assert params[i].dtype == np.int64
    params[i] = np.asarray(7, dtype=np.int64)
    m.set_parameters(params)
    assert int(m.get_model().bn1.num_batches_tracked) == 7
    assert sum(p.size for p in resnet18_model(seed=0).get_parameters() if p.dtype == np.float32) > 11_000_000


def test_dataset_partitions_and_exports():
    data = P2PFLDataset.synthetic_mnist(1003, 200)
    assert data.get_num_samples(True) == 1003 and data.get(0)["image"].shape == (28, 28)
    for strat, kw in [(RandomIIDPartitionStrategy, {}), (DirichletPartitionStrategy, {"alpha": 0.5}), (LabelSkewedPartitionStrategy, {}), (PercentageBasedNonIIDPartitionStrategy, {"iid_fraction": 0.3})]:
        parts = data.generate_partitions(7, strat, **kw)
        assert sum(p.get_num_samples(True) for p in parts) == 1003
        assert sum(p.get_num_samples(False) for p in parts) == 200
    parts = data.generate_partitions(4, RandomIIDPartitionStrategy)
    sizes = [p.get_num_samples(True) for p in parts]
    assert max(sizes) - min(sizes) <= 1
    x, y = parts[0].export(TensorExportStrategy, train=True)
    assert x.shape == (sizes[0], 28, 28) and y.dtype == torch.int64
    loader = parts[0].export(PyTorchExportStrategy, train=True, batch_size=16)
    batch = next(iter(loader))
    assert batch["image"].shape == (16, 28, 28) and batch["label"].shape == (16,)


def test_dirichlet_huge_alpha_is_near_uniform_and_small_alpha_is_skewed():
    data = P2PFLDataset.synthetic_mnist(4000, 400)
    even = data.generate_partitions(4, DirichletPartitionStrategy, alpha=1e6, min_partition_size=10)
    for p in even:
        labels = np.asarray(p.get(slice(None))["label"])
        counts = np.bincount(labels, minlength=10) / len(labels)
        assert np.all(np.abs(counts - 0.1) < 0.05)
    skew = data.generate_partitions(4, DirichletPartitionStrategy, alpha=0.05, min_partition_size=1)
    frac = [np.bincount(np.asarray(p.get(slice(None))["label"]), minlength=10).max() / p.get_num_samples() for p in skew]
    assert max(frac) > 0.3
    with pytest.raises(ValueError):
        data.generate_partitions(4, DirichletPartitionStrategy, alpha=-1.0)


def test_unsplit_dataset_and_split():
    d = P2PFLDataset.from_arrays({"image": np.zeros((50, 4)), "label": np.arange(50) % 5})
    with pytest.raises(ValueError):
        d.get_num_samples()
    d.generate_train_test_split(test_size=0.2, seed=1)
    assert d.get_num_samples(True) == 40 and d.get_num_samples(False) == 10
    import pandas as pd

    dp = P2PFLDataset.from_pandas(pd.DataFrame({"image": [1.0, 2.0, 3.0, 4.0], "label": [0, 1, 0, 1]}))
    dp.generate_train_test_split(test_size=0.5)
    assert dp.get_num_samples(True) == 2
    dg = P2PFLDataset.from_generator(lambda: ({"image": [i], "label": i % 2} for i in range(10)))
    dg.generate_train_test_split(test_size=0.3)
    assert dg.get_num_samples(False) == 3


def test_topology_matrices():
    f = TopologyFactory.generate_matrix
    np.testing.assert_array_equal(f(TopologyType.STAR, 4), [[0, 1, 1, 1], [1, 0, 0, 0], [1, 0, 0, 0], [1, 0, 0, 0]])
    np.testing.assert_array_equal(f(TopologyType.FULL, 4), [[0, 1, 1, 1], [1, 0, 1, 1], [1, 1, 0, 1], [1, 1, 1, 0]])
    np.testing.assert_array_equal(f(TopologyType.LINE, 4), [[0, 1, 0, 0], [1, 0, 1, 0], [0, 1, 0, 1], [0, 0, 1, 0]])
    np.testing.assert_array_equal(f(TopologyType.RING, 4), [[0, 1, 0, 1], [1, 0, 1, 0], [0, 1, 0, 1], [1, 0, 1, 0]])
    with pytest.raises(TypeError):
        f(3, 4)
    for seed in range(5):
        m = f(TopologyType.RANDOM, 12, p=0.1, seed=seed)
        assert (m == m.T).all() and m.diagonal().sum() == 0
        reach = np.linalg.matrix_power(m + np.eye(12, dtype=int), 12)
        assert (reach > 0).all()  # connected
    assert (f("random", 9, seed=1) == f("random", 9, seed=1)).all()


def test_connect_nodes_call_pattern():
    from unittest.mock import MagicMock

    nodes = [MagicMock(addr=f"n{i}") for i in range(4)]
    TopologyFactory.connect_nodes(TopologyFactory.generate_matrix(TopologyType.LINE, 4), nodes)
    nodes[0].connect.assert_called_once_with("n1")
    nodes[1].connect.assert_called_once_with("n2")
    nodes[2].connect.assert_called_once_with("n3")
    nodes[3].connect.assert_not_called()


def test_checkpoint_roundtrip(tmp_path):
    m = mlp_model(seed=5)
    m.set_contribution(["a", "b"], 77)
    path = str(tmp_path / "ck" / "model.ckpt")
    save_checkpoint(path, m, round=3, total_rounds=10, exp_name="experiment")
    m2 = mlp_model(seed=6)
    meta = load_checkpoint(path, m2)
    assert meta["round"] == 3 and meta["total_rounds"] == 10 and m2.get_num_samples() == 77
    for a, b in zip(m.get_parameters(), m2.get_parameters()):
        np.testing.assert_array_equal(a, b)
    with pytest.raises(ModelNotMatchingError):
        load_checkpoint(path, cnn_model(seed=0))


def _write_idx(path, arr):
    import struct

    arr = np.ascontiguousarray(arr, dtype=np.uint8)
    with open(path, "wb") as f:
        f.write(struct.pack(">HBB", 0, 8, arr.ndim))
        for d in arr.shape:
            f.write(struct.pack(">I", d))
        f.write(arr.tobytes())


def test_torchvision_importer_writes_parquet_and_a_card_that_load_back(tmp_path, monkeypatch):
    """Reference role: torchvision_to_datasets.py:31-190 (dataset -> Hugging Face layout -> Hub).  A tiny fabricated MNIST on disk goes
    through the CLI (no download, no network): parquet shards with the Hub's Image struct + a dataset card; both our loader and the
    in-memory path read the same pixels back.  Pushing needs a network: the CLI reports that instead of raising."""
    pytest.importorskip("torchvision")
    pytest.importorskip("pyarrow")
    from myfyp_b200.learning.dataset import torchvision_import as tvi

    rng = np.random.default_rng(0)
    raw = tmp_path / "MNIST" / "raw"
    raw.mkdir(parents=True)
    xs = {"train": rng.integers(0, 256, (20, 28, 28), dtype=np.uint8), "t10k": rng.integers(0, 256, (8, 28, 28), dtype=np.uint8)}
    ys = {"train": rng.integers(0, 10, 20, dtype=np.uint8), "t10k": rng.integers(0, 10, 8, dtype=np.uint8)}
    for k in xs:
        _write_idx(raw / f"{k}-images-idx3-ubyte", xs[k])
        _write_idx(raw / f"{k}-labels-idx1-ubyte", ys[k])
    out = tmp_path / "hf"
    rc = tvi.main(["--dataset_name", "MNIST", "--cache_dir", str(tmp_path), "--out_dir", str(out), "--no_download", "--dataset_summary", "tiny", "--official_link", "http://example.org"])
    assert rc == 0 and (out / "train.parquet").exists() and (out / "test.parquet").exists()
    card = (out / "README.md").read_text()
    assert "train: 20 examples" in card and "test: 8 examples" in card and "http://example.org" in card and "0 - zero" in card
    import pyarrow.parquet as pq

    t = pq.read_table(out / "train.parquet")
    assert t.num_rows == 20 and t.column_names == ["image", "label"] and b"huggingface" in t.schema.metadata
    from PIL import Image
    import io

    first = np.asarray(Image.open(io.BytesIO(t.column("image")[0].as_py()["bytes"])))
    np.testing.assert_array_equal(first, xs["train"][0])
    assert t.column("label").to_pylist() == [int(v) for v in ys["train"]]
    mem = tvi.from_torchvision("MNIST", str(tmp_path), download=False)
    assert mem.get_num_samples(train=True) == 20 and mem.get_num_samples(train=False) == 8
    np.testing.assert_allclose(np.asarray(mem.get(0, train=True)["image"]), xs["train"][0].astype(np.float32) / 255.0)
    # a push that cannot reach the Hub fails with a plain message and exit code 2, after the files were written
    def unreachable(self, *a, **k):
        raise ConnectionError("no route to huggingface.co")

    import huggingface_hub

    monkeypatch.setattr(huggingface_hub.HfApi, "create_repo", unreachable)
    rc = tvi.main(["--dataset_name", "MNIST", "--cache_dir", str(tmp_path), "--out_dir", str(out), "--no_download", "--token", "hf_x", "--repo_id", "nobody/tiny-mnist"])
    assert rc == 2
