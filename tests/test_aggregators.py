"""Aggregator math and lifecycle (modelled on the reference's test/learning/aggregator_test.py and scaffold_test.py)."""

import numpy as np
import pytest
import torch

from myfyp_b200.learning.aggregators import FedAvg, FedMedian, FedProx, Krum, NoModelsToAggregateError, Scaffold, TrimmedMean
from myfyp_b200.learning.p2pfl_model import P2PFLModel
from myfyp_b200.models import mlp_model


class ListModel(P2PFLModel):
    """Framework-free model: a list of arrays (the reference tests use the same kind of mock)."""

    def __init__(self, model=None, params=None, num_samples=None, contributors=None, additional_info=None):
        self._params = []
        super().__init__(model, params, num_samples, contributors, additional_info)

    def get_parameters(self):
        return self._params

    def set_parameters(self, params):
        self._params = [np.asarray(p) for p in params]

    def build_copy(self, **kw):
        return ListModel(None, **kw)

    def get_framework(self):
        return "mock"


def test_fedavg_toy_arrays_weighted():
    a = ListModel(params=[np.array([1.0, 2.0]), np.array([[1.0]])], num_samples=1, contributors=["a"])
    b = ListModel(params=[np.array([3.0, 6.0]), np.array([[5.0]])], num_samples=3, contributors=["b"])
    out = FedAvg().aggregate([a, b])
    np.testing.assert_allclose(out.get_parameters()[0], [2.5, 5.0])
    np.testing.assert_allclose(out.get_parameters()[1], [[4.0]])
    assert out.get_contributors() == ["a", "b"] and out.get_num_samples() == 4


def test_fedavg_slab_models_plus_minus_one():
    """theta, theta+1 (x2 samples), theta-1 (x2 samples) -> theta   (reference aggregator_test.py:83-113, atol 1e-7)."""
    base = mlp_model(seed=0)
    ref = base.get_parameters()
    m1 = base.build_copy(params=ref, num_samples=1, contributors=["1"])
    m2 = base.build_copy(params=[p + 1 for p in ref], num_samples=2, contributors=["2"])
    m3 = base.build_copy(params=[p - 1 for p in ref], num_samples=2, contributors=["3"])
    out = FedAvg().aggregate([m1, m2, m3])
    for a, b in zip(out.get_parameters(), ref):
        np.testing.assert_allclose(a, b, atol=1e-6)
    assert set(out.get_contributors()) == {"1", "2", "3"} and out.get_num_samples() == 5


def test_aggregator_lifecycle_and_partial_aggregation():
    agg = FedAvg("n")
    agg.set_nodes_to_aggregate(["a", "b", "c"])
    with pytest.raises(Exception):
        agg.set_nodes_to_aggregate(["x"])  # aggregation running
    ma = ListModel(params=[np.array([1.0])], num_samples=1, contributors=["a"])
    mb = ListModel(params=[np.array([3.0])], num_samples=1, contributors=["b"])
    assert agg.add_model(ma) == ["a"]
    assert agg.add_model(ma) == []  # duplicate contributor
    assert agg.add_model(ListModel(params=[np.array([0.0])], num_samples=1, contributors=["zzz"])) == []  # not in train set
    assert agg.add_model(ListModel(params=[np.array([0.0])], num_samples=1)) == []  # no contributors (reference bug Q4)
    assert set(agg.add_model(mb)) == {"a", "b"}
    assert agg.get_missing_models() == {"c"}
    part = agg.get_model(except_nodes=["a"])  # what a peer that already holds `a` still needs
    assert part.get_contributors() == ["b"]
    both = agg.get_model(except_nodes=[])
    np.testing.assert_allclose(both.get_parameters()[0], [2.0])
    with pytest.raises(NoModelsToAggregateError):
        agg.get_model(except_nodes=["a", "b"])
    out = agg.wait_and_get_aggregation(timeout=0.2)  # times out (c missing) and aggregates what it has
    np.testing.assert_allclose(out.get_parameters()[0], [2.0])
    agg.clear()
    agg.set_nodes_to_aggregate(["a"])


def test_non_partial_aggregators_offer_raw_models():
    for agg in (FedMedian(), Krum(f=0), Scaffold()):
        assert agg.partial_aggregation is False
    agg = FedMedian("n")
    agg.set_nodes_to_aggregate(["a", "b"])
    ma = ListModel(params=[np.array([1.0])], num_samples=1, contributors=["a"])
    agg.add_model(ma)
    assert agg.get_model(except_nodes=[]) is ma


@pytest.mark.parametrize("k", [1, 2, 3, 4, 5, 8])
def test_fedmedian_matches_numpy(k):
    rng = np.random.default_rng(k)
    base = mlp_model(seed=0)
    plist = [[rng.standard_normal(p.shape).astype(np.float32) for p in base.get_parameters()] for _ in range(k)]
    models = [base.build_copy(params=p, num_samples=1, contributors=[str(i)]) for i, p in enumerate(plist)]
    out = FedMedian().aggregate(models)
    for i, got in enumerate(out.get_parameters()):
        np.testing.assert_allclose(got, np.median(np.array([p[i] for p in plist]), axis=0), atol=1e-6)
    assert len(out.get_contributors()) == k
    gen = FedMedian().aggregate([ListModel(params=p, num_samples=1, contributors=[str(i)]) for i, p in enumerate(plist)])
    for a, b in zip(gen.get_parameters(), out.get_parameters()):
        np.testing.assert_allclose(a, b, atol=1e-6)


def test_trimmed_mean_drops_outliers():
    base = mlp_model(seed=0)
    ref = base.get_parameters()
    vals = [0.0, 1.0, 2.0, 3.0, 1000.0]
    models = [base.build_copy(params=[p * 0 + v for p in ref], num_samples=1, contributors=[str(i)]) for i, v in enumerate(vals)]
    out = TrimmedMean(trim=1).aggregate(models)
    np.testing.assert_allclose(out.get_parameters()[0], np.full_like(ref[0], 2.0), atol=1e-6)


def test_krum_picks_the_cluster_and_ignores_byzantine():
    torch.manual_seed(0)
    base = mlp_model(seed=0)
    ref = base.get_parameters()
    rng = np.random.default_rng(0)
    good = [[p + 0.01 * rng.standard_normal(p.shape).astype(np.float32) for p in ref] for _ in range(5)]
    bad = [[-10 * p for p in ref], [p + 5 for p in ref]]
    models = [base.build_copy(params=p, num_samples=10, contributors=[f"n{i}"]) for i, p in enumerate(good + bad)]
    k = Krum(f=2, m=1)
    out = k.aggregate(models)
    assert k.last_selected[0] < 5
    np.testing.assert_allclose(out.get_parameters()[0], good[k.last_selected[0]][0], atol=1e-6)
    assert len(out.get_contributors()) == 7  # the result stands for the whole train set
    mk = Krum(f=2, m=3)
    mk.aggregate(models)
    assert all(i < 5 for i in mk.last_selected)
    # generic (non-slab) path agrees on the selection
    kg = Krum(f=2, m=1)
    kg.aggregate([ListModel(params=p, num_samples=10, contributors=[f"n{i}"]) for i, p in enumerate(good + bad)])
    assert kg.last_selected == k.last_selected


@pytest.mark.parametrize("poison", ["nan", "inf", "huge"])
@pytest.mark.parametrize("where", [0, 3])
def test_krum_survives_non_finite_models(poison, where):
    """The simplest Byzantine update: a model full of NaN / Inf / 1e30.  It must rank last wherever it sits in the list (index 0
    is the model the Gram matrix is centred on), the honest ranking must match the NumPy oracle, and nothing may be NaN."""
    from myfyp_b200 import ops
    from myfyp_b200.learning.aggregators.krum import krum_select_reference

    torch.manual_seed(1)
    base = torch.randn(4096)
    slabs = [base + 0.01 * torch.randn(4096) for _ in range(6)]
    slabs[where] = torch.full((4096,), {"nan": float("nan"), "inf": float("inf"), "huge": 1e30}[poison])
    dist = ops.pairwise_sqdist(slabs)
    assert not torch.isnan(dist).any()
    others = [i for i in range(6) if i != where]
    assert torch.isinf(dist[where, others]).all() and torch.isinf(dist[others, where]).all()
    assert torch.isfinite(dist[others][:, others]).all()
    want = torch.cdist(torch.stack([slabs[i] for i in others]).double(), torch.stack([slabs[i] for i in others]).double()) ** 2
    torch.testing.assert_close(dist[others][:, others].double(), want, rtol=1e-3, atol=1e-6)
    scores, sel = ops.krum_select(dist, 1, 5)
    assert where not in sel and sorted(sel) == others
    assert (scores, sel) == krum_select_reference(dist.numpy(), 1, 5)
    # more broken models than f allows for: the honest ones still rank first
    slabs[(where + 1) % 6] = torch.full((4096,), float("nan"))
    _, sel2 = ops.krum_select(ops.pairwise_sqdist(slabs), 0, 4)
    assert where not in sel2 and (where + 1) % 6 not in sel2


def _scaffold_inputs():
    # the reference's own fixture values (test/learning/scaffold_test.py:80-141)
    p = [np.array([1.0, 2.0]), np.array([3.0])]
    infos = [
        {"delta_y_i": [np.array([0.1, 0.2]), np.array([0.3])], "delta_c_i": [np.array([0.01, 0.02]), np.array([0.03])]},
        {"delta_y_i": [np.array([0.3, 0.2]), np.array([0.1])], "delta_c_i": [np.array([0.03, 0.02]), np.array([0.01])]},
    ]
    return p, infos, [10, 30]


def test_scaffold_exact_values_generic():
    p, infos, ns = _scaffold_inputs()
    models = [ListModel(params=p, num_samples=n, contributors=[f"n{i}"], additional_info={"scaffold": info}) for i, (info, n) in enumerate(zip(infos, ns))]
    agg = Scaffold(global_lr=0.1)
    out = agg.aggregate(models)
    # x = p + 0.1 * (10*dy0 + 30*dy1)/40 ; c = (dc0 + dc1)/2
    np.testing.assert_allclose(out.get_parameters()[0], [1.0 + 0.1 * (0.1 * 10 + 0.3 * 30) / 40, 2.0 + 0.1 * 0.2], atol=1e-7)
    np.testing.assert_allclose(out.get_parameters()[1], [3.0 + 0.1 * (0.3 * 10 + 0.1 * 30) / 40], atol=1e-7)
    c = out.get_info("scaffold")["global_c"]
    np.testing.assert_allclose(c[0], [0.02, 0.02], atol=1e-7)
    np.testing.assert_allclose(c[1], [0.02], atol=1e-7)
    out2 = agg.aggregate(models)  # second round accumulates on the kept global state
    np.testing.assert_allclose(out2.get_info("scaffold")["global_c"][0], [0.04, 0.04], atol=1e-7)


def test_scaffold_slab_path_matches_generic():
    base = mlp_model(seed=0)
    ref = base.get_parameters()
    rng = np.random.default_rng(1)
    infos = [{"delta_y_i": [rng.standard_normal(p.shape).astype(np.float32) for p in ref], "delta_c_i": [rng.standard_normal(p.shape).astype(np.float32) for p in ref]} for _ in range(3)]
    ns = [5, 10, 20]
    slab_models = [base.build_copy(params=ref, num_samples=n, contributors=[str(i)], additional_info={"scaffold": info}) for i, (info, n) in enumerate(zip(infos, ns))]
    gen_models = [ListModel(params=ref, num_samples=n, contributors=[str(i)], additional_info={"scaffold": info}) for i, (info, n) in enumerate(zip(infos, ns))]
    # rebase=False: the reference's math verbatim (x starts as models[0]'s parameters and persists in the aggregator)
    a, b = Scaffold(global_lr=0.5, rebase=False).aggregate(slab_models), Scaffold(global_lr=0.5).aggregate(gen_models)
    for x, y in zip(a.get_parameters(), b.get_parameters()):
        np.testing.assert_allclose(x, y, atol=1e-5)
    ca = a.get_info("scaffold")["global_c"]
    lay = base.layout
    for e, y in zip(lay.in_order(), b.get_info("scaffold")["global_c"]):
        np.testing.assert_allclose(ca[e.offset : e.offset + e.numel].view(e.shape).numpy(), y, atol=1e-5)


def test_scaffold_rebases_on_the_round_start_recovered_from_the_clients():
    """Default slab path: x = y_k - delta_y_k, so an aggregator that sat out rounds (node outside the train set) does not apply
    this round's deltas to a stale copy of the global model."""
    base = mlp_model(seed=0)
    x = [p.copy() for p in base.get_parameters()]
    rng = np.random.default_rng(2)
    dys = [[0.01 * rng.standard_normal(p.shape).astype(np.float32) for p in x] for _ in range(2)]
    ns = [10, 30]
    models = [
        base.build_copy(params=[p + d for p, d in zip(x, dy)], num_samples=n, contributors=[str(i)], additional_info={"scaffold": {"delta_y_i": dy, "delta_c_i": [np.zeros_like(p) for p in x]}})
        for i, (dy, n) in enumerate(zip(dys, ns))
    ]
    agg = Scaffold(global_lr=1.0)
    agg.global_model_params = torch.full((base.layout.total,), 123.0)  # a stale leftover from some earlier round
    out = agg.aggregate(models)
    for got, p, d0, d1 in zip(out.get_parameters(), x, dys[0], dys[1]):
        np.testing.assert_allclose(got, p + 0.25 * d0 + 0.75 * d1, atol=1e-5)


def test_scaffold_survives_every_transport_form():
    """Deltas that travelled as bytes (per-layer lists, the reference's wire form), as a flat array (descriptor side channel)
    or stayed tensors must aggregate to the same model; a malformed shape raises instead of being broadcast over layers."""
    import pickle

    from myfyp_b200.learning.aggregators.scaffold import _as_flat
    from myfyp_b200.models import mlp_model

    torch.manual_seed(0)
    base = mlp_model(seed=0)
    clients = []
    for k in range(3):
        m = base.build_copy(params=base.slab.clone() + 0.01 * (k + 1), num_samples=10 * (k + 1), contributors=[f"n{k}"])
        m.add_info("scaffold", {"delta_y_i": torch.arange(m.layout.total, dtype=torch.float32) * 1e-4 * (k + 1), "delta_c_i": torch.full((m.layout.total,), 0.5 * (k + 1))})
        clients.append(m)
    want = Scaffold(global_lr=0.5).aggregate(clients)
    # (a) bytes round trip: encode -> build_copy(bytes) on the receiving side
    received = [base.build_copy(params=m.encode_parameters(), num_samples=m.get_num_samples(), contributors=m.get_contributors()) for m in clients]
    wire = pickle.loads(clients[0].encode_parameters())["additional_info"]["scaffold"]["delta_y_i"]
    assert isinstance(wire, list) and len(wire) == len(base.layout.order)
    got = Scaffold(global_lr=0.5).aggregate(received)

    def same(a, b):  # compare the state_dict entries (slab padding is not part of the model and does not travel)
        for x, y in zip(a, b):
            np.testing.assert_allclose(x, y, rtol=1e-6, atol=1e-6)

    from myfyp_b200.parallel.slab import layers_from_flat

    same(got.get_parameters(), want.get_parameters())
    same(layers_from_flat(_as_flat(got.get_info("scaffold")["global_c"], base), base.layout), layers_from_flat(want.get_info("scaffold")["global_c"], base.layout))
    # (b) flat array form
    flat = [base.build_copy(params=m.slab.clone(), num_samples=m.get_num_samples(), contributors=m.get_contributors()) for m in clients]
    for f, m in zip(flat, clients):
        f.add_info("scaffold", {k: v.numpy() for k, v in m.get_info("scaffold").items()})
    same(Scaffold(global_lr=0.5).aggregate(flat).get_parameters(), want.get_parameters())
    rt = _as_flat(np.arange(base.layout.total, dtype=np.float32), base)
    torch.testing.assert_close(rt, torch.arange(base.layout.total, dtype=torch.float32))
    # (c) malformed
    with pytest.raises(ValueError):
        _as_flat(np.zeros(7, dtype=np.float32), base)
    with pytest.raises(ValueError):
        _as_flat([np.zeros(3)], base)


def test_scaffold_errors():
    with pytest.raises(NoModelsToAggregateError):
        Scaffold().aggregate([])
    m = ListModel(params=[np.array([1.0])], num_samples=1, contributors=["a"], additional_info={"scaffold": {"delta_y_i": [np.array([1.0])]}})
    with pytest.raises(ValueError):
        Scaffold().aggregate([m])
    assert Scaffold().get_required_callbacks() == ["scaffold"]
    assert FedProx(mu=0.1).get_required_callbacks() == ["fedprox"]


def test_empty_aggregation_raises():
    for agg in (FedAvg(), FedMedian(), Krum()):
        with pytest.raises(NoModelsToAggregateError):
            agg.aggregate([])


def test_coordinate_shards_cover_the_model_once_and_keep_16_byte_alignment():
    """ops._shard_bounds: contiguous, aligned, complete - for any model size and worker count (the GPU side is tested on 2 GPUs)."""
    from myfyp_b200 import ops

    for n in (1, 1023, 1024, 262144, 262145, 11_173_962, (1 << 20) + 1539):
        for parts in (1, 2, 3, 7, 8):
            b = ops._shard_bounds(n, parts)
            assert len(b) == parts + 1 and b[0] == 0 and b[-1] == n
            assert all(x <= y for x, y in zip(b, b[1:]))
            assert all(lo % 1024 == 0 for lo, hi in zip(b, b[1:]) if hi > lo)  # every non-empty slice starts 4 KB-aligned
    # on the CPU (or for a small model) a reduction is never sharded
    slabs = [torch.zeros(1 << 19) for _ in range(3)]
    assert ops.shard_devices(slabs, torch.device("cpu")) == [torch.device("cpu")]
    assert ops.shard_devices(slabs, torch.device("cpu"), force=True) == [torch.device("cpu")]
