"""Whole federations on the GPU(s) of the box: every robust aggregator must keep the honest nodes learning under the config-3
attack (sign-flip attackers, random topology, train set smaller than the network), Scaffold and FedProx must learn with a train
set smaller than the network.  Runs once on one GPU and once spread over every visible GPU (peer access inside one process) -
the second case is where aggregates used to be born on the wrong GPU (FedMedian 0.10, Scaffold 0.13 in round 1).
Reference behaviour: p2pfl/learning/aggregators/scaffold.py:70-122; FedMedian's intended math fedmedian.py:55-65; Krum, TrimmedMean
and FedProx do not exist in the reference."""

import os
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

N_GPUS = torch.cuda.device_count() if torch.cuda.is_available() else 0
GPU_CASES = [1] + ([N_GPUS] if N_GPUS > 1 else [])


def _agg(name):
    from myfyp_b200.learning.aggregators import FedMedian, FedProx, Krum, Scaffold, TrimmedMean

    return {
        "fedmedian": FedMedian, "trimmedmean": lambda: TrimmedMean(trim=2), "krum": lambda: Krum(f=2, m=2),
        "scaffold": lambda: Scaffold(global_lr=1.0, variate_scaling="paper"), "fedprox": lambda: FedProx(mu=0.01),
    }[name]


@pytest.mark.skipif(N_GPUS == 0, reason="needs a GPU")
@pytest.mark.parametrize("gpus", GPU_CASES)
@pytest.mark.parametrize("agg", ["fedmedian", "trimmedmean", "krum"])
def test_robust_aggregators_keep_honest_nodes_learning_under_attack(agg, gpus):
    from benchmarks.configs import run

    out = run(f"t-{agg}-{gpus}", 12, "cnn", _agg(agg), "random", 8, 512, 6, "cuda", byzantine=2, max_gpus=gpus)
    assert out["gpus"] == gpus
    assert out["honest_acc_min"] is not None and out["honest_acc_min"] > 0.9, out


@pytest.mark.skipif(N_GPUS == 0, reason="needs a GPU")
@pytest.mark.parametrize("gpus", GPU_CASES)
@pytest.mark.parametrize("agg", ["scaffold", "fedprox"])
def test_scaffold_and_fedprox_learn_with_a_partial_train_set(agg, gpus):
    from benchmarks.configs import run

    # Scaffold's control variates are SGD arithmetic ((x - y) / (K lr) estimates a gradient only under SGD): its clients run SGD
    # (1024 samples per node and 8 rounds = 128 plain-SGD steps: with 512 samples and 6 rounds the least-elected node sat at the 0.9
    # line and the test failed about one run in four; lr 0.1 diverges on this CNN)
    kw = {"optimizer": "sgd", "lr": 0.05} if agg == "scaffold" else None
    out = run(f"t-{agg}-{gpus}", 12, "cnn", _agg(agg), "random", 8, 1024, 8, "cuda", max_gpus=gpus, learner_kw=kw)
    assert out["honest_acc_min"] is not None and out["honest_acc_min"] > 0.9, out


@pytest.mark.skipif(N_GPUS == 0, reason="needs a GPU")
def test_arena_refuses_a_descriptor_whose_slot_was_reused():
    """Slots of the publish arena rotate.  A descriptor is valid for the generation it was issued for: resolving it after the owner
    has published something newer into the same slot must raise (the command layer then refuses the model) instead of silently
    handing out the newer bytes.  Publication itself does not synchronise the host."""
    from myfyp_b200.learning.exceptions import StaleDescriptorError
    from myfyp_b200.parallel.ipc import ArenaPool

    dev = torch.device("cuda", 0)
    pool = ArenaPool(0, 1, dev, slot_elems=1 << 20, slots=2)
    try:
        a, b, c = (torch.full((1 << 20,), float(v), device=dev) for v in (1, 2, 3))
        sa = pool.publish(a)
        assert float(pool.view(sa.rank, sa.offset, 1 << 20, sa.generation)[123]) == 1.0
        sb = pool.publish(b)
        sc = pool.publish(c)  # wraps around: reuses a's slot
        assert sc.offset == sa.offset and sc.generation != sa.generation
        with pytest.raises(StaleDescriptorError):
            pool.view(sa.rank, sa.offset, 1 << 20, sa.generation)
        assert not pool.still_valid(sa.rank, sa.offset, sa.generation) and pool.still_valid(sb.rank, sb.offset, sb.generation)
        assert float(pool.view(sb.rank, sb.offset, 1 << 20, sb.generation)[5]) == 2.0
        assert float(pool.view(sc.rank, sc.offset, 1 << 20, sc.generation)[5]) == 3.0
        assert pool.stale_refused == 1
    finally:
        pool.close()


@pytest.mark.skipif(N_GPUS < 2, reason="NVLS needs at least two GPUs")
def test_nvls_in_switch_fedavg_matches_fp64():
    """multimem.ld_reduce / multimem.st all-reduce over every visible GPU (one process, one communicator per GPU) against an fp64
    weighted average; two calls in a row (epoch handling), odd sizes."""
    from myfyp_b200.parallel.collective import NvlsFedAvg, nvls_available

    if not nvls_available():
        pytest.skip("the box does not expose NVSwitch multicast to this process")
    devs = [torch.device("cuda", i) for i in range(N_GPUS)]
    for n in (1 << 20, 4 * 12345):
        comms = NvlsFedAvg.in_process(n, devs)
        try:
            for trial in range(2):
                g = torch.Generator().manual_seed(n + trial)
                xs = [torch.randn(n, generator=g) for _ in devs]
                ws = torch.rand(len(devs), generator=g)
                ws = (ws / ws.sum()).tolist()
                for c, x, w in zip(comms, xs, ws):
                    with torch.cuda.device(c.device):
                        c.src.copy_((x * w).to(c.device))
                for c in comms:
                    with torch.cuda.device(c.device):
                        c()
                for c in comms:
                    torch.cuda.synchronize(c.device)
                    assert not c.timed_out()
                want = sum(x.double() * w for x, w in zip(xs, ws)).float()
                for c in comms:
                    torch.testing.assert_close(c.out.cpu(), want, rtol=1e-5, atol=1e-6)
        finally:
            for c in comms:
                c.close()
