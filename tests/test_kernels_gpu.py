"""Numerics of every hand-written kernel against a plain PyTorch fp32 (or fp64) reference of the same op."""

import math

import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dev():
    from myfyp_b200.ops import native

    native.load()  # fail loudly if the library is missing on a GPU box
    return torch.device("cuda", 0)


def _slabs(k, n, dev, seed=0, scale=1.0):
    g = torch.Generator(device="cpu").manual_seed(seed)
    return [(torch.randn(n, generator=g) * scale).to(dev) for _ in range(k)]


@pytest.mark.parametrize("k,n", [(1, 64), (2, 4096), (5, 235264), (8, 1 << 20), (64, 65536), (70, 8192), (3, 1027)])
def test_fedavg(dev, k, n):
    from myfyp_b200 import ops

    slabs = _slabs(k, n, dev)
    w = torch.rand(k).tolist()
    out = ops.fedavg(slabs, w)
    ref = sum(s.double() * wi for s, wi in zip(slabs, w)).float()
    torch.testing.assert_close(out, ref, rtol=1e-5, atol=1e-5)


def test_fedavg_unaligned_views(dev):
    from myfyp_b200 import ops

    base = _slabs(3, 5001, dev)
    slabs = [b[1:] for b in base]  # 4-byte aligned only -> scalar path
    out = ops.fedavg(slabs, [0.2, 0.3, 0.5])
    ref = 0.2 * slabs[0] + 0.3 * slabs[1] + 0.5 * slabs[2]
    torch.testing.assert_close(out, ref, rtol=1e-5, atol=1e-6)


def test_accumulate_and_pull_copy(dev):
    from myfyp_b200 import ops

    a, b = _slabs(2, 100003, dev)
    acc = a.clone()
    ops.accumulate(acc, b, 0.25, beta=0.5)
    torch.testing.assert_close(acc, 0.5 * a + 0.25 * b, rtol=1e-6, atol=1e-6)
    dst = torch.empty_like(a)
    ops.pull_copy(dst, b)
    assert torch.equal(dst, b)


@pytest.mark.parametrize("k", [1, 2, 3, 4, 7, 8, 16, 31, 32, 33, 64])
def test_median(dev, k):
    from myfyp_b200 import ops

    slabs = _slabs(k, 40000, dev, seed=k)
    out = ops.median(slabs)
    st = torch.stack(slabs).sort(dim=0).values
    ref = (st[(k - 1) // 2] + st[k // 2]) / 2  # NumPy's median (what the reference intends, fedmedian.py:59-62)
    torch.testing.assert_close(out, ref, rtol=1e-6, atol=1e-6)


@pytest.mark.parametrize("k,trim", [(5, 1), (8, 2), (16, 3), (40, 10)])
def test_trimmed_mean(dev, k, trim):
    from myfyp_b200 import ops

    slabs = _slabs(k, 30000, dev, seed=k)
    out = ops.median(slabs, trim=trim)
    st = torch.stack(slabs).sort(dim=0).values
    torch.testing.assert_close(out, st[trim : k - trim].mean(dim=0), rtol=1e-5, atol=1e-6)




def test_krum_gpu_path_survives_non_finite_models(dev):
    """NaN / 1e30 model at index 0 (the centre of the Gram matrix) and at index > 0: distances to it are +inf, never NaN, the
    selection equals the NumPy oracle's and `sel` holds valid indices only."""
    import numpy as np

    from myfyp_b200 import ops
    from myfyp_b200.learning.aggregators.krum import krum_select_reference

    torch.manual_seed(2)
    base = torch.randn(1 << 16, device=dev)
    for where in (0, 4):
        for poison in (float("nan"), 1e30):
            slabs = [base + 0.01 * torch.randn_like(base) for _ in range(7)]
            slabs[where] = torch.full_like(base, poison)
            dist = ops.pairwise_sqdist(slabs)
            assert not torch.isnan(dist).any()
            others = [i for i in range(7) if i != where]
            assert torch.isinf(dist[where, others]).all()
            scores, sel = ops.krum_select(dist, 1, 6)
            ref_scores, ref_sel = krum_select_reference(dist.cpu().numpy(), 1, 6)
            assert sel == ref_sel and where not in sel
            np.testing.assert_allclose(scores, ref_scores, rtol=1e-5)


@pytest.mark.parametrize("k,n,passes", [(2, 4096, 3), (5, 235264, 3), (16, 1 << 20, 3), (64, 131072, 3), (33, 70000, 3), (8, 1 << 20, 1)])
def test_krum_gram_tcgen05(dev, k, n, passes):
    """Pairwise distances from the tcgen05 Gram kernel vs fp64 direct distances; models close to each other (late-round regime)."""
    from myfyp_b200 import ops

    g = torch.Generator().manual_seed(k)
    base = torch.randn(n, generator=g)
    slabs = [(base + 0.01 * torch.randn(n, generator=g) * (1 + i % 3)).to(dev) for i in range(k)]
    d = ops.pairwise_sqdist(slabs, passes=passes).double().cpu()
    m = torch.stack([s.double().cpu() for s in slabs])
    ref = torch.cdist(m, m).pow(2)
    tol = 1e-4 if passes == 3 else 5e-3
    torch.testing.assert_close(d, ref, rtol=tol, atol=tol * ref.max().item())
    # what matters for Krum: the ranking
    from myfyp_b200.learning.aggregators.krum import krum_select_reference

    f = max(0, (k - 3) // 2)
    s_ref, sel_ref = krum_select_reference(ref.numpy(), f, 1)
    s_dev, sel_dev = ops.krum_select(ops.pairwise_sqdist(slabs, passes=passes), f, 1)
    if passes == 3:
        assert sel_dev == sel_ref
        torch.testing.assert_close(torch.tensor(s_dev, dtype=torch.float64), torch.tensor(s_ref, dtype=torch.float64), rtol=1e-3, atol=1e-6)


def test_krum_gram_twice_same_workspace(dev):
    from myfyp_b200 import ops

    slabs = _slabs(6, 50000, dev)
    a = ops.pairwise_sqdist(slabs)
    b = ops.pairwise_sqdist(slabs)
    assert torch.equal(a, b)  # deterministic reduction + self-resetting counter


def test_krum_rejects_byzantine(dev):
    from myfyp_b200 import ops

    good = _slabs(6, 20000, dev, scale=0.01)
    bad = [(-5.0 * good[0]).contiguous(), (good[1] + 3.0).contiguous()]
    scores, order = ops.krum_select(ops.pairwise_sqdist(good + bad), 2, 8)
    assert set(order[-2:]) == {6, 7}


def test_scaffold_kernels(dev):
    from myfyp_b200 import ops

    n, k = 50048, 5
    x, c = _slabs(2, n, dev, seed=1)
    dys, dcs = _slabs(k, n, dev, seed=2), _slabs(k, n, dev, seed=3)
    wy = torch.rand(k).tolist()
    xr = x.double() + sum(w * d.double() for w, d in zip(wy, dys))
    cr = c.double() + sum(d.double() for d in dcs) / k
    ops.scaffold_server_update(x, c, dys, dcs, wy, 1.0 / k)
    torch.testing.assert_close(x, xr.float(), rtol=1e-5, atol=1e-5)
    torch.testing.assert_close(c, cr.float(), rtol=1e-5, atol=1e-5)
    ci, xs, y = _slabs(3, n, dev, seed=4)
    ci0 = ci.clone()
    dy, dc = torch.empty(n, device=dev), torch.empty(n, device=dev)
    ops.scaffold_finalize(ci, dy, dc, xs, y, 19 * 1e-3)
    adj = (xs - y) / (19 * 1e-3)
    torch.testing.assert_close(dc, adj, rtol=1e-5, atol=1e-5)
    torch.testing.assert_close(ci, ci0 + adj, rtol=1e-5, atol=1e-4)
    torch.testing.assert_close(dy, y - xs)


@pytest.mark.parametrize("terms", ["plain", "prox", "scaffold", "all"])
def test_adam_matches_torch(dev, terms):
    from myfyp_b200 import ops

    n = 70016
    p0, c, ci, ref_p = _slabs(4, n, dev, seed=5)
    p = p0.clone()
    pt = p0.clone().requires_grad_(True)
    opt = torch.optim.Adam([pt], lr=1e-3)
    m, v = torch.zeros(n, device=dev), torch.zeros(n, device=dev)
    kw = {}
    if terms in ("prox", "all"):
        kw.update(prox_mu=0.1, prox_ref=ref_p)
    if terms in ("scaffold", "all"):
        kw.update(cv_global=c, cv_local=ci, cv_scale=1.0)
    for step in range(1, 4):
        g = _slabs(1, n, dev, seed=10 + step)[0]
        eff = g.clone()
        if "prox_ref" in kw:
            eff += 0.1 * (pt.detach() - ref_p)
        if "cv_global" in kw:
            eff += c - ci
        pt.grad = eff
        opt.step()
        gbuf = g.clone()
        ops.adam_step(p, gbuf, m, v, step, lr=1e-3, **kw)
        assert float(gbuf.abs().max()) == 0.0  # gradient slab cleared by the kernel
    torch.testing.assert_close(p, pt.detach(), rtol=1e-4, atol=1e-5)


def test_sgd_momentum_matches_torch(dev):
    from myfyp_b200 import ops

    n = 30016
    p0 = _slabs(1, n, dev)[0]
    p, pt = p0.clone(), p0.clone().requires_grad_(True)
    opt = torch.optim.SGD([pt], lr=0.05, momentum=0.9, weight_decay=1e-3)
    buf = torch.zeros(n, device=dev)
    for step in range(3):
        g = _slabs(1, n, dev, seed=step + 1)[0]
        pt.grad = g.clone()
        opt.step()
        ops.sgd_step(p, g.clone(), buf, lr=0.05, momentum=0.9, weight_decay=1e-3)
    torch.testing.assert_close(p, pt.detach(), rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("b,c", [(1, 10), (32, 10), (257, 10), (64, 100), (16, 1000)])
def test_xent(dev, b, c):
    from myfyp_b200 import ops

    logits = (torch.randn(b, c, device=dev) * 3).requires_grad_(True)
    labels = torch.randint(0, c, (b,), device=dev)
    loss = ops.cross_entropy(logits, labels)
    loss.backward()
    ref_logits = logits.detach().clone().requires_grad_(True)
    ref = torch.nn.functional.cross_entropy(ref_logits, labels)
    ref.backward()
    torch.testing.assert_close(loss, ref, rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(logits.grad, ref_logits.grad, rtol=1e-3, atol=1e-6)
    acc = torch.zeros(2, device=dev)
    ops.cross_entropy_eval(logits.detach(), labels, acc)
    assert abs(acc[0].item() - ref.item() * b) < 1e-3 * max(1.0, ref.item() * b)
    assert acc[1].item() == (logits.argmax(1) == labels).sum().item()


def _bf(t):
    return t.to(torch.bfloat16)


@pytest.mark.parametrize("m,n,k", [(128, 64, 64), (32, 256, 784), (100, 10, 128), (256, 784, 32), (10, 128, 24), (300, 130, 200), (1024, 256, 784)])
def test_gemm_tcgen05_plain(dev, m, n, k):
    """D = A B^T on tcgen05/TMEM fed by TMA vs fp32 matmul of the same bf16-rounded operands."""
    from myfyp_b200 import ops

    pad8 = lambda x: (x + 7) // 8 * 8  # noqa: E731
    a = torch.zeros(m, pad8(k), dtype=torch.bfloat16, device=dev)
    b = torch.zeros(n, pad8(k), dtype=torch.bfloat16, device=dev)
    a[:, :k] = _bf(torch.randn(m, k, device=dev))
    b[:, :k] = _bf(torch.randn(n, k, device=dev))
    out = torch.full((m, n), float("nan"), device=dev)
    ops.gemm_bf16(a, b, m, n, k, out_f32=out)
    ref = a[:, :k].float() @ b[:, :k].float().T
    torch.testing.assert_close(out, ref, rtol=1e-3, atol=1e-3 * math.sqrt(k))


def test_gemm_tcgen05_epilogues(dev):
    from myfyp_b200 import ops

    m, n, k = 96, 136, 264
    a, b = _bf(torch.randn(m, k, device=dev)), _bf(torch.randn(n, k, device=dev))
    bias = torch.randn(n, device=dev)
    mask = _bf(torch.randn(m, n, device=dev))
    ref = a.float() @ b.float().T + bias
    # bias + relu -> bf16 and transposed bf16
    o16 = torch.zeros(m, n, dtype=torch.bfloat16, device=dev)
    o16t = torch.zeros(n, 96, dtype=torch.bfloat16, device=dev)
    ops.gemm_bf16(a, b, m, n, k, out_bf16=o16, out_bf16_t=o16t, bias=bias, relu=True)
    r = torch.relu(ref)
    torch.testing.assert_close(o16.float(), r, rtol=2e-2, atol=2e-2)
    torch.testing.assert_close(o16t.float(), r.T, rtol=2e-2, atol=2e-2)
    # mask + column sums + fp32 accumulate
    of = torch.ones(m, n, device=dev)
    cs = torch.zeros(n, device=dev)
    ops.gemm_bf16(a, b, m, n, k, out_f32=of, mask=mask, colsum=cs, accumulate=True)
    r2 = (a.float() @ b.float().T) * (mask.float() > 0)
    torch.testing.assert_close(of, r2 + 1.0, rtol=1e-3, atol=2e-2)
    torch.testing.assert_close(cs, r2.sum(0), rtol=1e-3, atol=5e-2)


def test_cast_transpose(dev):
    from myfyp_b200 import ops

    r, c = 77, 203
    src = torch.randn(r, c, device=dev)
    d = torch.zeros(r, 208, dtype=torch.bfloat16, device=dev)
    dt = torch.zeros(c, 80, dtype=torch.bfloat16, device=dev)
    cs = torch.zeros(c, device=dev)
    ops.cast_transpose(src, r, c, dst=d, dst_t=dt, colsum=cs)
    assert torch.equal(d[:, :c], src.to(torch.bfloat16))
    assert torch.equal(dt[:, :r], src.to(torch.bfloat16).T)
    torch.testing.assert_close(cs, src.sum(0), rtol=1e-5, atol=1e-4)


def _mlp_bf16_emulation(module, x, y):
    """The same forward/backward in plain PyTorch fp32 math, with bf16 rounding exactly where the kernels round."""
    bf = lambda t: t.to(torch.bfloat16).float()  # noqa: E731
    batch = x.shape[0]
    W = [bf(l.weight.detach()) for l in module.layers]
    b = [l.bias.detach() for l in module.layers]
    L = len(W)
    h = [bf(x.reshape(batch, -1).float())]
    for i in range(L):
        z = h[i] @ W[i].T + b[i]
        if i < L - 1:
            h.append(bf(torch.relu(z)))
    logits = z
    loss = torch.nn.functional.cross_entropy(logits, y)
    dl = torch.softmax(logits, 1)
    dl[torch.arange(batch), y] -= 1
    dl /= batch
    gw, gb = [None] * L, [None] * L
    dz, gb[L - 1] = bf(dl), dl.sum(0)
    for i in range(L - 1, -1, -1):
        gw[i] = dz.T @ h[i]
        if i > 0:
            dzf = (dz @ W[i]) * (h[i] > 0)
            gb[i - 1] = dzf.sum(0)
            dz = bf(dzf)
    return loss, gw, gb


@pytest.mark.parametrize("batch", [32, 24, 256])
def test_native_mlp_step_matches_reference(dev, batch):
    """tcgen05 forward/backward of the MLP: tight vs the fp32 PyTorch restatement with bf16-rounded operands,
    loose vs true-fp32 autograd (bf16 operand noise under the cancellation of a random-init gradient is 3-6 %)."""
    from myfyp_b200.models import mlp_model

    model = mlp_model(device="cuda:0", seed=3, native=True)
    module = model.get_model()
    x = torch.randn(batch, 28, 28, device=dev)
    y = torch.randint(0, 10, (batch,), device=dev)
    loss = torch.zeros(1, device=dev)
    module.native.refresh_weights()
    module.native_train_step(x, y, loss)
    grads = model.grad_slab.clone()
    model.grad_slab.zero_()
    ref_loss = torch.nn.functional.cross_entropy(module(x), y)
    ref_loss.backward()
    ref = model.grad_slab.clone()
    model.grad_slab.zero_()
    emu_loss, gw, gb = _mlp_bf16_emulation(module, x, y)
    assert abs(loss.item() - emu_loss.item()) < 1e-4 * max(1.0, abs(emu_loss.item()))
    assert abs(loss.item() - ref_loss.item()) < 2e-2 * max(1.0, abs(ref_loss.item()))
    for i in range(len(gw)):
        for name, g in ((f"layers.{i}.weight", gw[i]), (f"layers.{i}.bias", gb[i])):
            e = model.layout.entries[name]
            a = grads[e.offset : e.offset + e.numel]
            assert (a - g.reshape(-1)).norm() <= 1e-3 * g.norm() + 1e-7, name
    assert (grads - ref).norm() / ref.norm() < 1e-1


def test_native_mlp_learns(dev):
    from myfyp_b200.learning.dataset import P2PFLDataset
    from myfyp_b200.learning.torch_learner import TorchLearner
    from myfyp_b200.models import mlp_model
    from myfyp_b200.ops import native

    data = P2PFLDataset.synthetic_mnist(2048, 512, seed=1)
    model = mlp_model(device="cuda:0", seed=0, native=True)
    learner = TorchLearner(model, data, "n0", batch_size=64)
    learner.set_epochs(2)
    before = learner.evaluate()["test_metric"]
    n0 = native.launches()
    learner.fit()
    assert native.launches() > n0
    after = learner.evaluate()
    assert after["test_metric"] > max(0.9, before)


def test_cnn_learner_trains_and_gradients_live_in_the_slab(dev):
    from myfyp_b200.learning.dataset import P2PFLDataset
    from myfyp_b200.learning.torch_learner import TorchLearner
    from myfyp_b200.models import cnn_model

    data = P2PFLDataset.synthetic_cifar10(1024, 256, seed=2)
    model = cnn_model(device="cuda:0", seed=0, native=False)
    learner = TorchLearner(model, data, "n0", batch_size=64)
    learner.set_epochs(2)
    before = learner.evaluate()["test_metric"]
    learner.fit()
    after = learner.evaluate()
    assert after["test_metric"] > max(0.8, before)
    lo, hi = model.grad_slab.data_ptr(), model.grad_slab.data_ptr() + 4 * model.grad_slab.numel()
    for p in model.get_model().parameters():
        assert p.grad is not None and lo <= p.grad.data_ptr() < hi  # autograd accumulated in place, into the gradient slab
    assert float(model.grad_slab.abs().max()) == 0.0  # and the fused optimizer cleared it


@pytest.mark.parametrize("kind", ["mlp", "cnn"])
def test_cuda_graph_epoch_equals_eager_epoch(dev, kind):
    """A fit replayed from the captured epoch graph ends at the same parameters as the same fit run kernel by kernel."""
    from myfyp_b200.learning.dataset import P2PFLDataset
    from myfyp_b200.learning.torch_learner import TorchLearner
    from myfyp_b200.models import cnn_model, mlp_model

    mk = (lambda: mlp_model(device="cuda:0", seed=5, native=True)) if kind == "mlp" else (lambda: cnn_model(device="cuda:0", seed=5, native=False))
    data = P2PFLDataset.synthetic_mnist(300, 64, seed=3) if kind == "mlp" else P2PFLDataset.synthetic_cifar10(300, 64, seed=3)
    outs = []
    for graph in (False, True):
        model = mk()
        learner = TorchLearner(model, data, "n", batch_size=64, shuffle=False, use_cuda_graph=graph)
        for _ in range(2):  # second fit = a pure replay
            learner.fit()
        outs.append((model.slab.clone(), learner.evaluate()))
    if kind == "mlp":  # our kernels are deterministic: replay == eager
        torch.testing.assert_close(outs[0][0], outs[1][0], rtol=2e-3, atol=2e-4)
    else:  # the autograd conv path may pick different (atomic, tf32) cuDNN algorithms under capture
        assert (outs[0][0] - outs[1][0]).norm() < 5e-2 * outs[0][0].norm()
    assert abs(outs[0][1]["test_loss"] - outs[1][1]["test_loss"]) < 5e-2


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_fedavg_pulls_from_peer_gpus():
    from myfyp_b200 import ops
    from myfyp_b200.parallel.fabric import Fabric

    Fabric.get().enable_peer_access()
    n = 1 << 20
    slabs = [torch.randn(n, device=f"cuda:{i % torch.cuda.device_count()}") for i in range(4)]
    out = ops.fedavg(slabs, [0.25] * 4, device=torch.device("cuda", 0))
    ref = sum(s.to("cuda:0") for s in slabs) * 0.25
    torch.testing.assert_close(out, ref, rtol=1e-5, atol=1e-5)
    med = ops.median(slabs, device=torch.device("cuda", 1))
    st = torch.stack([s.to("cuda:1") for s in slabs]).sort(dim=0).values
    torch.testing.assert_close(med, (st[1] + st[2]) / 2, rtol=1e-6, atol=1e-6)


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_median_and_krum_gram_are_sharded_by_coordinate_over_the_gpus_that_hold_the_models():
    """with contributors on two GPUs each GPU reduces one slice of the coordinates (peer reads, peer stores into
    the destination); the result equals the single-launch one bit for bit (median) / to fp32 summation order (Gram)."""
    from myfyp_b200 import ops

    n = (1 << 20) + 1536 + 3  # not a multiple of the shard alignment
    g = torch.Generator().manual_seed(3)
    slabs = [torch.randn(n, generator=g).to(f"cuda:{i % 2}") for i in range(7)]
    dev0 = torch.device("cuda", 0)
    assert [d.index for d in ops.shard_devices(slabs, dev0, force=True)] == [1, 0]
    assert [d.index for d in ops.shard_devices(slabs, dev0)] == [0]  # automatic: from three GPUs on
    before = ops.native.launches()
    sharded = ops.median(slabs, device=dev0, shard=True)
    assert ops.native.launches() - before == 2  # one launch per GPU
    single = ops.median(slabs, device=dev0, shard=False)
    torch.cuda.synchronize()
    assert torch.equal(sharded, single)
    ref = torch.stack([s.to(dev0) for s in slabs]).median(dim=0).values
    torch.testing.assert_close(sharded, ref, rtol=0, atol=0)
    tm = ops.median(slabs, trim=2, device=torch.device("cuda", 1), shard=True)
    st = torch.stack([s.to("cuda:1") for s in slabs]).sort(dim=0).values
    torch.testing.assert_close(tm, st[2:5].mean(dim=0), rtol=1e-5, atol=1e-6)
    slabs = [s[: n - 3] for s in slabs]  # the Gram kernel takes whole 16-byte groups (slab layouts guarantee them)
    dist = ops.pairwise_sqdist(slabs, device=dev0, shard=True)
    big = torch.stack([s.to(dev0).double() for s in slabs])
    want = torch.cdist(big, big) ** 2
    torch.testing.assert_close(dist.double(), want, rtol=1e-4, atol=1e-2)
    assert [int(i) for i in dist.sum(1).argsort()] == [int(i) for i in want.sum(1).argsort()]


@pytest.mark.parametrize("world,n", [(2, 4096), (4, 235264), (8, 1 << 20), (3, 1000 * 4)])
def test_twoshot_fedavg_emulated_ranks(dev, world, n):
    """The fused reduce-scatter + all-gather kernel with all ranks in one launch (the single-GPU way to run ranks that wait on each other)."""
    from myfyp_b200.parallel.collective import TwoShotFedAvg

    coll = TwoShotFedAvg.emulated(n, world, dev)
    g = torch.Generator().manual_seed(world)
    for call in range(3):  # epochs advance, flags and counters are reused
        xs = [torch.randn(n, generator=g).to(dev) for _ in range(world)]
        w = torch.rand(world, generator=g)
        w = (w / w.sum()).tolist()
        for r in range(world):
            coll.srcs[r].copy_(xs[r])
            coll.outs[r].fill_(float("nan"))
        coll(w)
        torch.cuda.synchronize(dev)
        ref = sum(x.double() * wi for x, wi in zip(xs, w)).float()
        for r in range(world):
            torch.testing.assert_close(coll.outs[r], ref, rtol=1e-5, atol=1e-5, msg=f"rank {r} call {call}")


@pytest.mark.parametrize("a_mn,b_mn", [(True, False), (False, True), (True, True)])
@pytest.mark.parametrize("m,n,k", [(128, 64, 64), (256, 784, 32), (10, 128, 24), (300, 130, 200), (64, 27, 4096)])
def test_gemm_tcgen05_mn_major_operands(dev, a_mn, b_mn, m, n, k):
    """Operands stored transposed ([K, M] / [K, N]) consumed as MN-major UMMA operands: D = A B^T must not change."""
    from myfyp_b200 import ops

    pad8 = lambda x: (x + 7) // 8 * 8  # noqa: E731
    a = _bf(torch.randn(m, k, device=dev))
    b = _bf(torch.randn(n, k, device=dev))
    a_store = torch.zeros(k, pad8(m), dtype=torch.bfloat16, device=dev) if a_mn else torch.zeros(m, pad8(k), dtype=torch.bfloat16, device=dev)
    b_store = torch.zeros(k, pad8(n), dtype=torch.bfloat16, device=dev) if b_mn else torch.zeros(n, pad8(k), dtype=torch.bfloat16, device=dev)
    if a_mn:
        a_store[:, :m] = a.T
    else:
        a_store[:, :k] = a
    if b_mn:
        b_store[:, :n] = b.T
    else:
        b_store[:, :k] = b
    out = torch.full((m, n), float("nan"), device=dev)
    ops.gemm_bf16(a_store, b_store, m, n, k, out_f32=out, a_mn=a_mn, b_mn=b_mn)
    ref = a.float() @ b.float().T
    torch.testing.assert_close(out, ref, rtol=1e-3, atol=1e-3 * math.sqrt(k))


@pytest.mark.parametrize("splits", [2, 7, 64])
def test_gemm_tcgen05_split_k(dev, splits):
    from myfyp_b200 import ops

    m, n, k = 32, 27, 65536  # the wgrad shape of a first conv layer: tiny output, very long reduction
    a_store = _bf(torch.randn(k, m, device=dev) * 0.1)  # [K, M]
    b_store = torch.zeros(k, 32, dtype=torch.bfloat16, device=dev)
    b_store[:, :n] = _bf(torch.randn(k, n, device=dev) * 0.1)
    out = torch.zeros(m, n, device=dev)
    ops.gemm_bf16(a_store, b_store, m, n, k, out_f32=out, a_mn=True, b_mn=True, split_k=splits)
    ref = a_store.float().T @ b_store[:, :n].float()
    torch.testing.assert_close(out, ref, rtol=2e-3, atol=2e-2)


@pytest.mark.parametrize("B,C,H,k,stride,pad,first", [(4, 3, 32, 3, 1, 1, True), (3, 32, 16, 3, 1, 1, False), (2, 64, 8, 3, 2, 1, False), (5, 128, 4, 4, 1, 0, False), (2, 16, 8, 1, 2, 0, False), (2, 12, 8, 3, 1, 1, False)])
def test_im2col_col2im_match_unfold_fold(dev, B, C, H, k, stride, pad, first):
    from myfyp_b200 import ops

    F = torch.nn.functional
    x = torch.randn(B, C, H, H, device=dev)
    xq = x.to(torch.bfloat16).float()
    oh = (H + 2 * pad - k) // stride + 1
    kd = C * k * k
    ld = (kd + 7) // 8 * 8
    col = torch.zeros(B * oh * oh, ld, dtype=torch.bfloat16, device=dev)
    src = x.contiguous() if first else xq.permute(0, 2, 3, 1).contiguous().to(torch.bfloat16)
    ops.im2col(src, col, B, H, H, C, k, k, stride, pad, nchw_f32=first)
    # unfold orders a patch (c, kh, kw); the kernels use (kh, kw, c) so that a tap's channels are contiguous
    ref = F.unfold(xq, k, padding=pad, stride=stride).transpose(1, 2).reshape(B * oh * oh, C, k * k).transpose(1, 2).reshape(B * oh * oh, kd)
    assert torch.equal(col[:, :kd].float(), ref)
    # col2im == fold (sum of overlapping patches), NHWC out
    dcol = torch.zeros_like(col)
    dcol[:, :kd] = torch.randn(B * oh * oh, kd, device=dev).to(torch.bfloat16)
    dx = torch.empty(B, H, H, C, dtype=torch.bfloat16, device=dev)
    ops.col2im(dcol, dx, B, H, H, C, k, k, stride, pad)
    dcol_ckk = dcol[:, :kd].float().reshape(B, oh * oh, k * k, C).transpose(2, 3).reshape(B, oh * oh, kd)
    folded = F.fold(dcol_ckk.transpose(1, 2), (H, H), k, padding=pad, stride=stride)
    torch.testing.assert_close(dx.float().permute(0, 3, 1, 2), folded, rtol=2e-2, atol=2e-2)


def test_maxpool_and_colsum(dev):
    from myfyp_b200 import ops

    F = torch.nn.functional
    B, H, C = 3, 8, 32
    y = torch.relu(torch.randn(B, H, H, C, device=dev)).to(torch.bfloat16)
    p = torch.empty(B, H // 2, H // 2, C, dtype=torch.bfloat16, device=dev)
    idx = torch.empty(B, H // 2, H // 2, C, dtype=torch.uint8, device=dev)
    ops.maxpool2_fwd(y, p, idx, B, H, H, C)
    yt = y.float().permute(0, 3, 1, 2).requires_grad_(True)
    ref = F.max_pool2d(yt, 2)
    assert torch.equal(p.float().permute(0, 3, 1, 2), ref.detach())
    dp = torch.randn(B, H // 2, H // 2, C, device=dev).to(torch.bfloat16)
    dz = torch.empty_like(y)
    ops.maxpool2_relu_bwd(dp, idx, y, dz, B, H, H, C)
    ref.backward(dp.float().permute(0, 3, 1, 2))
    expect = (yt.grad * (yt.detach() > 0)).permute(0, 2, 3, 1)
    torch.testing.assert_close(dz.float(), expect, rtol=0, atol=0)
    out = torch.zeros(C, device=dev)
    ops.colsum_bf16(dz.view(-1, C), B * H * H, C, out)
    torch.testing.assert_close(out, dz.float().view(-1, C).sum(0), rtol=1e-4, atol=1e-4)


class _RoundFwd(torch.autograd.Function):
    """bf16 rounding in the forward pass only (where our kernels store an activation as bf16)."""

    @staticmethod
    def forward(ctx, x):
        return x.to(torch.bfloat16).float()

    @staticmethod
    def backward(ctx, g):
        return g


class _RoundBwd(torch.autograd.Function):
    """bf16 rounding of the gradient only (where our kernels store a gradient tensor as bf16)."""

    @staticmethod
    def forward(ctx, x):
        return x.view_as(x)

    @staticmethod
    def backward(ctx, g):
        return g.to(torch.bfloat16).float()


def _cnn_bf16_emulation(module, x, y):
    """The CNN step in plain PyTorch fp32 math with bf16 rounding where the hand-written path rounds (weights, stored
    activations, stored gradients); gradients w.r.t. the fp32 master weights by straight-through."""
    F = torch.nn.functional
    convs = [m for m in module.features if isinstance(m, torch.nn.Conv2d)]
    fcs = [m for m in module.classifier if isinstance(m, torch.nn.Linear)]
    params = [p for m in convs + fcs for p in (m.weight, m.bias)]
    w = lambda m: m.weight + (m.weight.to(torch.bfloat16).float() - m.weight).detach()  # noqa: E731  value = bf16(W), grad -> W
    a = _RoundFwd.apply(x.float())
    for m in convs:
        z = F.conv2d(_RoundBwd.apply(a), w(m), m.bias, stride=m.stride, padding=m.padding)
        a = F.max_pool2d(_RoundFwd.apply(torch.relu(_RoundBwd.apply(z))), 2)
    h = _RoundFwd.apply(torch.relu(_RoundBwd.apply(F.linear(_RoundBwd.apply(a.flatten(1)), w(fcs[0]), fcs[0].bias))))
    logits = _RoundBwd.apply(F.linear(h, w(fcs[1]), fcs[1].bias))
    loss = F.cross_entropy(logits, y)
    grads = torch.autograd.grad(loss, params)
    return loss.detach(), logits.detach(), dict(zip([n for m_ in ("features", "classifier") for n, _ in getattr(module, m_).named_parameters(prefix=m_)], grads))


@pytest.mark.parametrize("batch", [16, 64])
def test_native_cnn_step_matches_reference(dev, batch):
    """Hand-written CNN forward/backward (patch gather + tcgen05 GEMMs): tight against the fp32 PyTorch restatement that
    rounds to bf16 where the kernels do, loose against true-fp32 autograd / cuDNN."""
    from myfyp_b200.models import cnn_model

    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    model = cnn_model(device="cuda:0", seed=2, native=True, precision="bf16")
    module = model.get_model()
    x = torch.randn(batch, 3, 32, 32, device=dev)
    y = torch.randint(0, 10, (batch,), device=dev)
    loss = torch.zeros(1, device=dev)
    module.native.refresh_weights()
    module.native_train_step(x, y, loss)
    grads = model.grad_slab.clone()
    model.grad_slab.zero_()
    logits_native = module.native.logits(x)
    emu_loss, emu_logits, emu_grads = _cnn_bf16_emulation(module, x, y)
    model.grad_slab.zero_()
    torch.testing.assert_close(logits_native, emu_logits, rtol=2e-2, atol=2e-3)
    assert abs(loss.item() - emu_loss.item()) < 1e-3 * max(1.0, abs(emu_loss.item()))
    for e in model.layout.in_order():
        a, b = grads[e.offset : e.offset + e.numel], emu_grads[e.name].reshape(-1)
        rel = ((a - b).norm() / (b.norm() + 1e-12)).item()
        assert rel < 3e-2, (e.name, rel)
    # true fp32 autograd: direction of the whole gradient
    ref_loss = torch.nn.functional.cross_entropy(module(x), y)
    ref_loss.backward()
    ref = model.grad_slab.clone()
    model.grad_slab.zero_()
    assert abs(loss.item() - ref_loss.item()) < 2e-2 * max(1.0, abs(ref_loss.item()))
    assert torch.nn.functional.cosine_similarity(grads, ref, dim=0).item() > 0.98


def test_native_cnn_learns(dev):
    from myfyp_b200.learning.dataset import P2PFLDataset
    from myfyp_b200.learning.torch_learner import TorchLearner
    from myfyp_b200.models import cnn_model

    data = P2PFLDataset.synthetic_cifar10(2048, 512, seed=4)
    model = cnn_model(device="cuda:0", seed=1, native=True, precision="bf16")
    assert getattr(model.get_model(), "native_train_step", None) is not None
    learner = TorchLearner(model, data, "n0", batch_size=64)
    learner.set_epochs(2)
    before = learner.evaluate()["test_metric"]
    learner.fit()
    after = learner.evaluate()
    assert after["test_metric"] > max(0.9, before), (before, after)


def test_fused_pool_im2col_and_unpool_col2im_match_the_unfused_kernels(dev):
    from myfyp_b200 import ops

    B, H, C, k, pad = 3, 16, 32, 3, 1  # Y is [B,H,H,C]; the next conv works on the pooled [B,H/2,H/2,C] map
    PH = H // 2
    y = torch.relu(torch.randn(B, H, H, C, device=dev)).to(torch.bfloat16)
    # unfused: pool -> im2col
    p = torch.empty(B, PH, PH, C, dtype=torch.bfloat16, device=dev)
    idx = torch.empty(B, PH, PH, C, dtype=torch.uint8, device=dev)
    ops.maxpool2_fwd(y, p, idx, B, H, H, C)
    kd = C * k * k
    col = torch.zeros(B * PH * PH, kd, dtype=torch.bfloat16, device=dev)
    ops.im2col(p, col, B, PH, PH, C, k, k, 1, pad)
    # fused
    idx2 = torch.empty_like(idx)
    col2 = torch.zeros_like(col)
    ops.pool_im2col(y, idx2, col2, B, PH, PH, C, k, k, 1, pad)
    assert torch.equal(idx, idx2) and torch.equal(col, col2)
    # backward: unfused col2im -> unpool+mask -> colsum  vs fused
    dcol = torch.randn(B * PH * PH, kd, device=dev).to(torch.bfloat16)
    dp = torch.empty_like(p)
    ops.col2im(dcol, dp, B, PH, PH, C, k, k, 1, pad)
    dz = torch.empty_like(y)
    ops.maxpool2_relu_bwd(dp, idx, y, dz, B, H, H, C)
    db = torch.zeros(C, device=dev)
    ops.colsum_bf16(dz.view(-1, C), B * H * H, C, db)
    dz2 = torch.full_like(y, float("nan"))
    db2 = torch.zeros(C, device=dev)
    ops.unpool_col2im(dcol, idx, y, dz2, db2, B, PH, PH, C, k, k, 1, pad)
    assert torch.equal(dz, dz2)
    torch.testing.assert_close(db, db2, rtol=1e-4, atol=1e-4)


def test_xent_train_matches_xent_plus_cast(dev):
    from myfyp_b200 import ops

    b, c = 64, 10
    logits = torch.randn(b, c, device=dev) * 2
    labels = torch.randint(0, c, (b,), device=dev)
    loss = torch.full((1,), 123.0, device=dev)
    dl16 = torch.zeros(b, 16, dtype=torch.bfloat16, device=dev)
    db = torch.zeros(c, device=dev)
    ops.xent_train(logits, labels, loss, dl16, db, b, c)
    ref_logits = logits.clone().requires_grad_(True)
    ref = torch.nn.functional.cross_entropy(ref_logits, labels)
    ref.backward()
    torch.testing.assert_close(loss[0], ref.detach(), rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(dl16[:, :c].float(), ref_logits.grad, rtol=1e-2, atol=1e-4)
    torch.testing.assert_close(db, ref_logits.grad.sum(0), rtol=1e-3, atol=1e-5)


def test_adam_baked_step_with_device_base_and_bf16_mirror(dev):
    """Two 'epochs' of 3 steps with step = device base + baked offset must equal torch.optim.Adam's 6 steps; the mirror tracks p."""
    from myfyp_b200 import ops

    n = 4096
    p0 = _slabs(1, n, dev)[0]
    p, pt = p0.clone(), p0.clone().requires_grad_(True)
    opt = torch.optim.Adam([pt], lr=1e-3)
    m, v, st = torch.zeros(n, device=dev), torch.zeros(n, device=dev), torch.zeros(4, device=dev)
    p16 = torch.zeros(n, dtype=torch.bfloat16, device=dev)
    k = 0
    for epoch in range(2):
        for s in range(3):
            k += 1
            g = _slabs(1, n, dev, seed=100 + k)[0]
            pt.grad = g.clone()
            opt.step()
            ops.adam_step(p, g.clone(), m, v, 0, lr=1e-3, dev_state=st, step_offset=s + 1, p16=p16)
        ops.adam_advance(st, 3)
    assert st[0].item() == 6.0
    torch.testing.assert_close(p, pt.detach(), rtol=1e-4, atol=1e-6)
    assert torch.equal(p16, p.to(torch.bfloat16))
    full = torch.zeros(n, dtype=torch.bfloat16, device=dev)
    ops.cast_slab(p, full)
    assert torch.equal(full, p16)


def test_batchnorm_kernels_match_torch(dev):
    from myfyp_b200 import ops

    M, C = 4096, 64
    zf = torch.randn(M, C, device=dev) * 2 + 0.5
    z = zf.to(torch.bfloat16)
    res = torch.randn(M, C, device=dev).to(torch.bfloat16)
    gamma, beta = torch.rand(C, device=dev) + 0.5, torch.randn(C, device=dev)
    rm, rv, nb = torch.zeros(C, device=dev), torch.ones(C, device=dev), torch.zeros(1, device=dev)
    stats = torch.cat([z.float().sum(0), (z.float() ** 2).sum(0)])
    y = torch.empty_like(z)
    ops.bn_apply(z, res, y, stats, gamma, beta, rm, rv, nb, M, C, relu=True)
    zt = z.float().clone().requires_grad_(True)
    gt, bt = gamma.clone().requires_grad_(True), beta.clone().requires_grad_(True)
    trm, trv = torch.zeros(C, device=dev), torch.ones(C, device=dev)
    ref = torch.relu(torch.nn.functional.batch_norm(zt, trm, trv, gt, bt, training=True, momentum=0.1, eps=1e-5) + res.float())
    torch.testing.assert_close(y.float(), ref.detach(), rtol=2e-2, atol=2e-2)
    torch.testing.assert_close(rm, trm, rtol=1e-3, atol=1e-4)
    torch.testing.assert_close(rv, trv, rtol=1e-3, atol=1e-4)
    assert nb.item() == 1.0
    dy = torch.randn(M, C, device=dev).to(torch.bfloat16)
    ref.backward(dy.float())
    g, dz, red = torch.empty_like(z), torch.empty_like(z), torch.zeros(2 * C, device=dev)
    dgamma, dbeta = torch.zeros(C, device=dev), torch.zeros(C, device=dev)
    ops.bn_backward(dy, y, z, g, dz, red, stats, gamma, dgamma, dbeta, M, C, relu=True)
    mask = (y.float() > 0)
    torch.testing.assert_close(g.float(), dy.float() * mask, rtol=0, atol=0)
    torch.testing.assert_close(dz.float(), zt.grad, rtol=3e-2, atol=3e-3)
    torch.testing.assert_close(dgamma, gt.grad, rtol=2e-2, atol=5e-2)
    torch.testing.assert_close(dbeta, bt.grad, rtol=2e-2, atol=5e-2)
    # eval mode uses the running statistics
    y2 = torch.empty_like(z)
    ops.bn_apply(z, None, y2, None, gamma, beta, rm, rv, None, M, C, relu=False)
    ref2 = torch.nn.functional.batch_norm(z.float(), rm, rv, gamma, beta, training=False, eps=1e-5)
    torch.testing.assert_close(y2.float(), ref2, rtol=2e-2, atol=2e-2)


@pytest.mark.parametrize("B,H,cin,cout", [(4, 32, 64, 64), (3, 16, 128, 64), (5, 8, 64, 128), (6, 8, 256, 256), (11, 4, 128, 192), (8, 4, 512, 512)])
def test_implicit_gemm_conv_matches_conv2d(dev, B, H, cin, cout):
    """Implicit-GEMM 3x3 convolution (4-D TMA boxes as operand tiles): forward, input gradient and weight gradient against
    torch conv2d + autograd in fp32 on the same bf16-rounded operands."""
    from myfyp_b200 import ops

    torch.backends.cudnn.allow_tf32 = False
    g = torch.Generator(device="cpu").manual_seed(B * 1000 + H + cin)
    x = torch.randn(B, cin, H, H, generator=g).to(dev).bfloat16().float()
    w = (torch.randn(cout, cin, 3, 3, generator=g) / (3 * cin**0.5)).to(dev)
    dz = torch.randn(B, cout, H, H, generator=g).to(dev).bfloat16().float()
    extra = torch.randn(B, cin, H, H, generator=g).to(dev).bfloat16().float()
    assert ops.conv_gemm_supported(H, H, cin, cout, 3, 1, 1)
    # weights: slab-style fp32 [cout, cin, 9] -> bf16 [cout, 9, cin]
    w16 = torch.zeros(cout, 9 * cin, dtype=torch.bfloat16, device=dev)
    table = torch.tensor([0, 0, cout, cin, 9, 9 * cin], dtype=torch.int64, device=dev)
    ops.repack_conv_weights(w.reshape(-1), w16.reshape(-1), table, 1, w.numel())
    torch.testing.assert_close(w16.float().view(cout, 9, cin), w.bfloat16().float().view(cout, cin, 9).transpose(1, 2))
    wr = w.bfloat16().float().requires_grad_(True)
    xr = x.clone().requires_grad_(True)
    z_ref = torch.nn.functional.conv2d(xr, wr, padding=1)
    z_ref.backward(dz)
    M = B * H * H
    x_nhwc = x.permute(0, 2, 3, 1).contiguous().bfloat16()
    dz_nhwc = dz.permute(0, 2, 3, 1).contiguous().bfloat16()
    # forward (+ batch-norm statistics in the epilogue)
    z = torch.zeros(M, cout, dtype=torch.bfloat16, device=dev)
    cs, cq = torch.zeros(cout, device=dev), torch.zeros(cout, device=dev)
    ops.conv_gemm(ops.CONV_FWD, x_nhwc, B, H, H, cin, cout, 3, weights=w16, out_bf16=z, colsum=cs, colsumsq=cq)
    zr = z_ref.detach().permute(0, 2, 3, 1).reshape(M, cout)
    torch.testing.assert_close(z.float(), zr, rtol=2e-2, atol=2e-2)
    torch.testing.assert_close(cs, zr.sum(0), rtol=1e-2, atol=5e-2)
    torch.testing.assert_close(cq, (zr * zr).sum(0), rtol=1e-2, atol=5e-2)
    # input gradient (+ the gradient of a second consumer)
    dx = torch.zeros(M, cin, dtype=torch.bfloat16, device=dev)
    ops.conv_gemm(ops.CONV_DGRAD, dz_nhwc, B, H, H, cin, cout, 3, weights=w16, out_bf16=dx, add=extra.permute(0, 2, 3, 1).contiguous().bfloat16())
    dxr = (xr.grad + extra).permute(0, 2, 3, 1).reshape(M, cin)
    torch.testing.assert_close(dx.float(), dxr, rtol=2e-2, atol=3e-2)
    # weight gradient straight into PyTorch's layout, split-K over pixel blocks
    for split in (1, 4):
        dw = torch.zeros(cout, cin, 3, 3, device=dev)
        ops.conv_gemm(ops.CONV_WGRAD, x_nhwc, B, H, H, cin, cout, 3, dz=dz_nhwc, out_f32=dw, split_k=split)
        torch.testing.assert_close(dw, wr.grad, rtol=1e-2, atol=1e-2 * float(wr.grad.abs().max()))


def test_native_resnet18_step_matches_autograd(dev):
    """Hand-written ResNet-18 step vs torch autograd / cuDNN in fp32 on the same weights (bf16-operand tolerances)."""
    from myfyp_b200.models import resnet18_model

    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    model = resnet18_model(device="cuda:0", seed=1, native=True)
    module = model.get_model()
    batch = 32
    x = torch.randn(batch, 3, 32, 32, device=dev)
    y = torch.randint(0, 10, (batch,), device=dev)
    state0 = model.slab.clone()
    loss = torch.zeros(1, device=dev)
    module.native.refresh_weights()
    module.native_train_step(x, y, loss)
    grads = model.grad_slab.clone()
    model.grad_slab.zero_()
    running_native = model.slab.clone()
    model.slab.copy_(state0)  # undo the running-stat updates, then let torch do the same step
    module.train()
    ref_logits = module(x)
    ref_loss = torch.nn.functional.cross_entropy(ref_logits, y)
    ref_loss.backward()
    ref = model.grad_slab.clone()
    model.grad_slab.zero_()
    assert abs(loss.item() - ref_loss.item()) < 3e-2 * max(1.0, abs(ref_loss.item())), (loss.item(), ref_loss.item())
    running_torch = model.slab.clone()
    # A randomly initialised 20-layer BN net amplifies bf16 rounding (the BN backward subtracts two means),
    # so the yardstick is PyTorch's own bf16 autocast step on the same weights: the hand-written step has
    # to be at least as close to the fp32 gradient as that one is, layer by layer.
    model.slab.copy_(state0)
    with torch.autocast("cuda", dtype=torch.bfloat16):
        ac_loss = torch.nn.functional.cross_entropy(module(x).float(), y)
    ac_loss.backward()
    autocast = model.grad_slab.clone()
    model.grad_slab.zero_()
    model.slab.copy_(running_torch)
    cosine = torch.nn.functional.cosine_similarity
    cos, cos_ac = cosine(grads, ref, dim=0).item(), cosine(autocast, ref, dim=0).item()
    assert cos > 0.93 and cos > cos_ac - 0.02, (cos, cos_ac)
    for name in ("conv1.weight", "layers.0.conv1.weight", "layers.2.shortcut.0.weight", "layers.7.bn2.weight", "layers.4.bn1.bias", "fc.weight", "fc.bias"):
        e = model.layout.entries[name]
        sl = slice(e.offset, e.offset + e.numel)
        c, c_ac = cosine(grads[sl], ref[sl], dim=0).item(), cosine(autocast[sl], ref[sl], dim=0).item()
        assert c > 0.9 and c > c_ac - 0.03, (name, c, c_ac)
        ratio = (grads[sl].norm() / ref[sl].norm()).item()
        assert 0.85 < ratio < 1.15, (name, ratio)
    for name in ("fc.weight", "fc.bias", "layers.7.bn2.bias"):
        e = model.layout.entries[name]
        sl = slice(e.offset, e.offset + e.numel)
        assert cosine(grads[sl], ref[sl], dim=0).item() > 0.995, name
    # BatchNorm running statistics were updated like PyTorch's
    for name in ("bn1.running_mean", "layers.3.bn2.running_var", "layers.2.shortcut.1.running_mean"):
        e = model.layout.entries[name]
        torch.testing.assert_close(running_native[e.offset : e.offset + e.numel], model.slab[e.offset : e.offset + e.numel], rtol=5e-2, atol=5e-3)


def test_native_resnet18_learns_and_evaluates(dev):
    from myfyp_b200.learning.dataset import P2PFLDataset
    from myfyp_b200.learning.torch_learner import TorchLearner
    from myfyp_b200.models import resnet18_model

    data = P2PFLDataset.synthetic_cifar10(1024, 256, seed=6)
    model = resnet18_model(device="cuda:0", seed=0, native=True)
    learner = TorchLearner(model, data, "n0", batch_size=64, use_cuda_graph=False)
    learner.set_epochs(2)
    learner.fit()
    after = learner.evaluate()
    assert after["test_metric"] > 0.9, after
    assert int(model.get_model().bn1.num_batches_tracked) == 32  # 2 epochs x 16 steps, counted on the device
    # the eval-mode forward of the nn.Module (running statistics) agrees with the native evaluation
    module = model.get_model()
    module.eval()
    x, y = data.export(__import__("myfyp_b200.learning.dataset", fromlist=["TensorExportStrategy"]).TensorExportStrategy, train=False, device="cuda:0")
    with torch.no_grad():
        acc_ref = (module(x).argmax(1) == y).float().mean().item()
    assert abs(acc_ref - after["test_metric"]) < 0.05


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs in one process")
def test_in_process_federation_across_two_gpus_learns(dev):
    """Nodes of ONE process placed on two GPUs exchange descriptors of slabs that peers read over NVLink.  A snapshot or an
    aggregation result must be complete on its GPU before a descriptor leaves the producing thread (Fabric.make_visible):
    a consumer on the other GPU is ordered against nothing, and reading the allocation's old contents (last round's
    snapshot) silently stops the federation from learning while all nodes still agree on the same model."""
    import time

    from myfyp_b200.communication.protocols.memory.memory_communication_protocol import InMemoryCommunicationProtocol
    from myfyp_b200.learning.aggregators import FedAvg
    from myfyp_b200.learning.dataset import P2PFLDataset, RandomIIDPartitionStrategy
    from myfyp_b200.models import cnn_model
    from myfyp_b200.node import Node
    from myfyp_b200.parallel.fabric import Fabric
    from myfyp_b200.utils.topologies import TopologyFactory
    from myfyp_b200.utils.utils import set_fast_settings

    set_fast_settings(train_set_size=3)
    Fabric.get().enable_peer_access()
    n = 6
    data = P2PFLDataset.synthetic((3, 32, 32), 10, 256 * n, 64 * n, seed=0)
    parts = data.generate_partitions(n, RandomIIDPartitionStrategy, seed=0)
    nodes = [Node(cnn_model(device=f"cuda:{i % 2}", seed=i), parts[i], address=f"mg-{i}", aggregator=FedAvg(), protocol=InMemoryCommunicationProtocol, batch_size=64) for i in range(n)]
    try:
        for nd in nodes:
            nd.start()
        TopologyFactory.connect_nodes(TopologyFactory.generate_matrix("full", n), nodes)
        time.sleep(0.5)
        nodes[0].set_start_learning(rounds=8, epochs=1)
        deadline = time.time() + 120
        while not all(nd.learning_workflow.finished for nd in nodes):
            assert time.time() < deadline
            time.sleep(0.05)
        accs = [nd.learner.evaluate()["test_metric"] for nd in nodes]
        assert min(accs) > 0.5, accs  # chance is 0.1 - which is what the unordered cross-GPU reads produced
    finally:
        for nd in nodes:
            nd.stop()
