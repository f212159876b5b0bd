"""Node-batched TF32 training step (csrc/train_tf32.cu) against plain PyTorch fp32 autograd of the same ``nn.Module``.

Every kernel of the step is checked on its own output first (so a failure names the stage), then the whole step (loss,
gradients as seen by Adam, updated parameters), then 300 steps side by side with ``torch.optim.Adam`` (the reference's
optimizer, pytorch/lightning_model.py:181-183), then the learner-level cohort path with several nodes training at once.
"""

import threading

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

cuda = torch.cuda.is_available()


def _setup(groups=2, batch=64, n=200, seed=0):
    from myfyp_b200.models import cnn_model
    from myfyp_b200.models.grouped_cnn import CnnPlan

    dev = torch.device("cuda", 0)
    models = [cnn_model(device=dev, seed=seed + g, precision="tf32") for g in range(groups)]
    spec = models[0].get_model().grouped_spec
    plan = CnnPlan(spec, models[0].layout, dev, groups, batch)
    g = torch.Generator().manual_seed(seed)
    data = []
    for _ in range(groups):
        x = torch.randn(n, 3, 32, 32, generator=g).to(dev)
        y = torch.randint(0, 10, (n,), generator=g).to(dev)
        perm = torch.randperm(n, generator=g).to(dev)
        data.append((x, y, perm))
    return dev, models, spec, plan, data


def _slots(models, data, lr=1e-3, steps=8):
    slots = []
    for m, (x, y, perm) in zip(models, data):
        nt = m.layout.n_trainable
        slots.append(dict(slab=m.slab, m=torch.zeros(nt, device=m.device), v=torch.zeros(nt, device=m.device), adam_state=torch.zeros(4, device=m.device),
                          x=x, y=y, perm=perm, losses=torch.zeros(max(steps, 2), device=m.device), n_samples=x.shape[0], lr=lr))
    return slots


def _rel(a, b):
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-12))


def _ref_forward(module, xb):
    """fp32 activations of every stage (TF32 off: the reference of record is plain fp32)."""
    feats = list(module.features)
    acts, h = [], xb
    for i in range(0, len(feats), 3):
        z = F.conv2d(h, feats[i].weight, feats[i].bias, padding=1)
        h = F.max_pool2d(F.relu(z), 2)
        acts.append(h)
    return acts


@pytest.mark.skipif(not cuda, reason="needs a GPU")
def test_every_kernel_of_the_step_against_fp32_autograd():
    from myfyp_b200.models import grouped_cnn as gc

    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    try:
        dev, models, spec, plan, data = _setup()
        slots = _slots(models, data)
        plan.bind(slots)
        plan.refresh()
        B, step = 64, 1
        refs = []
        for m, (x, y, perm) in zip(models, data):
            mod = m.get_model()
            mod.zero_grad(set_to_none=True)
            idx = perm[step * B:(step + 1) * B]
            xb, yb = x[idx], y[idx]
            acts = _ref_forward(mod, xb)
            logits = mod(xb)
            loss = F.cross_entropy(logits, yb)
            loss.backward()
            refs.append((acts, logits.detach(), float(loss), xb, yb))
        # ---- forward, stage by stage
        plan.run(gc.RUN_CONV0_FWD, step)
        for g in range(2):
            got = plan.buffer(g, gc.BUF_ACT + 0).view(B, 16, 16, 32).permute(0, 3, 1, 2)
            assert _rel(got, refs[g][0][0].detach()) < 4e-3, f"conv0_fwd slot {g}"
        for l in (1, 2):
            plan.run(gc.RUN_CONV_FWD + l, step)
        for g in range(2):
            got = plan.buffer(g, gc.BUF_ACT + 1).view(B, 8, 8, 64).permute(0, 3, 1, 2)
            assert _rel(got, refs[g][0][1].detach()) < 6e-3, f"conv1_fwd slot {g}"
            got = (plan.buffer(g, gc.BUF_AHI) + plan.buffer(g, gc.BUF_ALO)).view(B, 4, 4, 128).permute(0, 3, 1, 2)
            assert _rel(got, refs[g][0][2].detach()) < 8e-3, f"conv2_fwd slot {g}"
        # fc1 (3xTF32) on OUR features: fp32-class agreement with an fp64 product of the same operands
        plan.run(gc.RUN_FC1_FWD, step)
        for g in range(2):
            mod = models[g].get_model()
            a = (plan.buffer(g, gc.BUF_AHI) + plan.buffer(g, gc.BUF_ALO)).view(B, 16, 128)
            w1 = mod.classifier[1].weight.view(256, 128, 16).permute(0, 2, 1).reshape(256, 2048)
            want = (a.reshape(B, 2048).double() @ w1.double().t()).float()
            got = plan.buffer(g, gc.BUF_Z1).view(B, 256)
            assert _rel(got, want) < 2e-5, f"fc1_fwd slot {g}: {_rel(got, want)}"
        plan.run(gc.RUN_HEAD_TRAIN, step)
        for g in range(2):
            assert abs(float(slots[g]["losses"][step]) - refs[g][2]) < 5e-3 * max(1.0, refs[g][2]), f"loss slot {g}"
        # ---- backward
        plan.run(gc.RUN_FC1_WGRAD, step)
        plan.run(gc.RUN_FC1_DGRAD, step)
        for l in (2, 1):
            plan.run(gc.RUN_CONV_WGRAD + l, step)
            plan.run(gc.RUN_CONV_DGRAD + l, step)
        plan.run(gc.RUN_CONV0_WGRAD, step)
        torch.cuda.synchronize()
        problems = []
        for g in range(2):
            mod, lay = models[g].get_model(), models[g].layout
            gm = plan.buffer(g, gc.BUF_GM)

            def grad_of(name, shape):
                e = lay.entries[name]
                return gm[e.offset:e.offset + e.numel].view(shape)

            checks = [
                ("classifier.3.weight", grad_of("classifier.3.weight", (10, 256)), mod.classifier[3].weight.grad, 2e-3),
                ("classifier.3.bias", grad_of("classifier.3.bias", (10,)), mod.classifier[3].bias.grad, 2e-3),
                ("classifier.1.bias", grad_of("classifier.1.bias", (256,)), mod.classifier[1].bias.grad, 1e-2),
                ("classifier.1.weight", grad_of("classifier.1.weight", (256, 16, 128)).permute(0, 2, 1).reshape(256, 2048), mod.classifier[1].weight.grad, 1e-2),
                ("features.6.weight", grad_of("features.6.weight", (128, 9, 64)).permute(0, 2, 1).reshape(128, 64, 3, 3), mod.features[6].weight.grad, 2e-2),
                ("features.6.bias", grad_of("features.6.bias", (128,)), mod.features[6].bias.grad, 2e-2),
                ("features.3.weight", grad_of("features.3.weight", (64, 9, 32)).permute(0, 2, 1).reshape(64, 32, 3, 3), mod.features[3].weight.grad, 2e-2),
                ("features.3.bias", grad_of("features.3.bias", (64,)), mod.features[3].bias.grad, 2e-2),
                ("features.0.weight", grad_of("features.0.weight", (32, 9, 3)).permute(0, 2, 1).reshape(32, 3, 3, 3), mod.features[0].weight.grad, 2e-2),
                ("features.0.bias", grad_of("features.0.bias", (32,)), mod.features[0].bias.grad, 2e-2),
            ]
            for name, got, want, tol in checks:
                r = _rel(got, want)
                cos = float(F.cosine_similarity(got.reshape(1, -1), want.reshape(1, -1)))
                if not (r < max(tol, 0.12) and cos > 0.998):  # TF32 vs fp32: a few ReLU / max-pool decisions flip; the exact-arithmetic test below is the tight one
                    problems.append(f"grad {name} slot {g}: rel {r:.3e} cos {cos:.6f}")
        assert not problems, "\n".join(problems)
    finally:
        torch.backends.cudnn.allow_tf32 = True


def _rna(t):
    """fp32 -> tf32, round to nearest (ties away), as cvt.rna.tf32.f32"""
    return ((t.contiguous().view(torch.int32) + 0x1000) & ~0x1FFF).view(torch.float32)


@pytest.mark.skipif(not cuda, reason="needs a GPU")
def test_backward_matches_an_fp32_restatement_with_the_same_rounding_points():
    """Index-exactness of every backward kernel: the same step restated in PyTorch fp32 with operands rounded to tf32 exactly where
    the kernels round them (products of tf32 numbers are exact in fp32, so only the summation order differs)."""
    from myfyp_b200.models import grouped_cnn as gc

    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    try:
        dev, models, spec, plan, data = _setup(groups=2, n=256, seed=5)
        slots = _slots(models, data)
        plan.bind(slots)
        plan.refresh()
        B, step = 64, 2
        for w in [gc.RUN_CONV0_FWD, gc.RUN_CONV_FWD + 1, gc.RUN_CONV_FWD + 2, gc.RUN_FC1_FWD, gc.RUN_HEAD_TRAIN, gc.RUN_FC1_WGRAD, gc.RUN_FC1_DGRAD,
                  gc.RUN_CONV_WGRAD + 2, gc.RUN_CONV_DGRAD + 2, gc.RUN_CONV_WGRAD + 1, gc.RUN_CONV_DGRAD + 1, gc.RUN_CONV0_WGRAD]:
            plan.run(w, step)
        torch.cuda.synchronize()
        problems = []
        with torch.no_grad():
            for g, (m, (x, y, perm)) in enumerate(zip(models, data)):
                mod, lay = m.get_model(), m.layout
                idx = perm[step * B:(step + 1) * B]
                xb, yb = _rna(x[idx]), y[idx]
                W = [_rna(mod.features[i].weight) for i in (0, 3, 6)]
                bs = [mod.features[i].bias for i in (0, 3, 6)]
                acts, pools, inds, h = [], [], [], xb
                for l in range(3):
                    z = F.conv2d(h, W[l], bs[l], padding=1)
                    # layers 0 and 1 store tf32-rounded activations and the kernels round BEFORE the 2x2 max (value and window
                    # position travel as one key); rounding is monotonic, so only which of two near-equal pixels wins can differ
                    p, ix = F.max_pool2d(_rna(F.relu(z)) if l < 2 else F.relu(z), 2, return_indices=True)
                    pools.append(p)
                    inds.append(ix)
                    h = _rna(p) if l < 2 else p
                    acts.append(h)
                W1, b1 = mod.classifier[1].weight, mod.classifier[1].bias
                W2, b2 = mod.classifier[3].weight, mod.classifier[3].bias
                feat = acts[2].reshape(B, -1)
                h1 = F.relu((feat.double() @ W1.double().t()).float() + b1)
                logits = h1 @ W2.t() + b2
                dl = (torch.softmax(logits, 1) - F.one_hot(yb, 10).float()) / B
                dh1 = (dl @ W2) * (h1 > 0)
                want = {"classifier.3.weight": dl.t() @ h1, "classifier.3.bias": dl.sum(0), "classifier.1.bias": dh1.sum(0),
                        "classifier.1.weight": (dh1.double().t() @ feat.double()).float()}
                da = (dh1.double() @ W1.double()).float().view(B, 128, 4, 4)
                xin = [xb, acts[0], acts[1]]
                for l in (2, 1, 0):
                    dm = _rna(da * (pools[l] > 0))
                    want[f"features.{3 * l}.bias"] = dm.sum((0, 2, 3))
                    dz = F.max_unpool2d(dm, inds[l], 2)
                    want[f"features.{3 * l}.weight"] = torch.nn.grad.conv2d_weight(xin[l], W[l].shape, dz, padding=1)
                    if l > 0:
                        da = torch.nn.grad.conv2d_input(xin[l].shape, W[l], dz, padding=1)
                gm = plan.buffer(g, gc.BUF_GM)
                for name, ref in want.items():
                    e = lay.entries[name]
                    got = gm[e.offset:e.offset + e.numel]
                    if name == "classifier.1.weight":
                        got = got.view(256, 16, 128).permute(0, 2, 1).reshape(256, 2048)
                    elif name.startswith("features") and name.endswith("weight"):
                        co, ci = ref.shape[0], ref.shape[1]
                        got = got.view(co, 9, ci).permute(0, 2, 1).reshape(co, ci, 3, 3)
                    else:
                        got = got.view(ref.shape)
                    scale = float(ref.abs().max())
                    # sums over 4 096 - 65 536 pixels of cancelling terms: the tensor core's fp32 accumulation order / rounding differs
                    # from cuDNN's, hence "off by more than 1 % of the largest entry" rather than bitwise
                    bad = float(((got - ref).abs() > 1e-2 * scale).float().mean())
                    cos = float(F.cosine_similarity(got.reshape(1, -1), ref.reshape(1, -1)))
                    if bad > 5e-3 or cos < 0.9999:
                        problems.append(f"{name} slot {g}: {bad:.2%} of the elements off by > 1 % of max, cos {cos:.7f}")
        assert not problems, "\n".join(problems)
    finally:
        torch.backends.cudnn.allow_tf32 = True


@pytest.mark.skipif(not cuda, reason="needs a GPU")
def test_full_step_updates_parameters_like_torch_adam_incl_partial_last_batch():
    dev, models, spec, plan, data = _setup(groups=2, n=200)
    torch.backends.cudnn.allow_tf32 = False
    try:
        import copy

        refs = [copy.deepcopy(m.get_model()) for m in models]
        opts = [torch.optim.Adam(r.parameters(), lr=1e-3) for r in refs]
        slots = _slots(models, data)
        plan.bind(slots)
        plan.refresh()
        for step in range(4):  # 200 samples: steps 0-2 full, step 3 has 8 samples
            n = plan.train_step(step)
            assert n == 13
            for g, (x, y, perm) in enumerate(data):
                idx = perm[step * 64:(step + 1) * 64]
                opts[g].zero_grad()
                loss = F.cross_entropy(refs[g](x[idx]), y[idx])
                loss.backward()
                opts[g].step()
                got = float(slots[g]["losses"][step])
                assert abs(got - float(loss)) < 1e-2 * max(1.0, float(loss)), f"step {step} slot {g}: loss {got} vs {float(loss)}"
        plan.advance(4)
        torch.cuda.synchronize()
        for g in range(2):
            assert float(slots[g]["adam_state"][0]) == 4.0
            for (name, p_ref), (_, p) in zip(refs[g].named_parameters(), models[g].get_model().named_parameters()):
                # Adam's first steps move every weight by ~lr whatever the gradient's size, so sign flips of tiny gradients are
                # the only source of difference: bound the mean, not the max
                d = (p.detach() - p_ref.detach()).abs()
                assert float(d.mean()) < 6e-4 and float(d.max()) < 8.1e-3, f"{name} slot {g}: mean {float(d.mean()):.2e} max {float(d.max()):.2e}"
    finally:
        torch.backends.cudnn.allow_tf32 = True


@pytest.mark.skipif(not cuda, reason="needs a GPU")
def test_300_steps_loss_curve_tracks_fp32_autograd():
    """The precision question of the benchmark: the native TF32 / 3xTF32 step and torch's fp32 autograd, same init, same
    batches, 300 Adam steps - the smoothed loss curves must stay within 10 % (+0.02) of each other all the way."""
    from myfyp_b200.learning.dataset import P2PFLDataset
    from myfyp_b200.learning.dataset.p2pfl_dataset import TensorExportStrategy
    from myfyp_b200.models import cnn_model
    from myfyp_b200.models.grouped_cnn import CnnPlan
    import copy

    dev = torch.device("cuda", 0)
    ds = P2PFLDataset.synthetic_cifar10(64 * 300, 64, seed=3)
    x, y = ds.export(TensorExportStrategy, train=True, device=dev)
    model = cnn_model(device=dev, seed=7, precision="tf32")
    ref = copy.deepcopy(model.get_model())
    opt = torch.optim.Adam(ref.parameters(), lr=1e-3)
    plan = CnnPlan(model.get_model().grouped_spec, model.layout, dev, 1, 64)
    nt = model.layout.n_trainable
    slot = dict(slab=model.slab, m=torch.zeros(nt, device=dev), v=torch.zeros(nt, device=dev), adam_state=torch.zeros(4, device=dev), x=x, y=y, perm=None,
                losses=torch.zeros(300, device=dev), n_samples=x.shape[0], lr=1e-3)
    plan.bind([slot])
    plan.refresh()
    ref_losses = []
    for s in range(300):
        plan.train_step(s)
        opt.zero_grad()
        loss = F.cross_entropy(ref(x[s * 64:(s + 1) * 64]), y[s * 64:(s + 1) * 64])
        loss.backward()
        opt.step()
        ref_losses.append(float(loss))
    ours = slot["losses"].cpu()
    theirs = torch.tensor(ref_losses)
    k = torch.ones(1, 1, 20) / 20
    so = F.conv1d(ours.view(1, 1, -1), k).view(-1)
    st = F.conv1d(theirs.view(1, 1, -1), k).view(-1)
    assert float(st[-1]) < 0.5 * float(st[0]), "the reference run itself must learn"
    gap = ((so - st).abs() - 0.1 * st - 0.02).max()
    assert float(gap) <= 0, f"loss curves diverge: worst excess {float(gap):.4f} (ours end {float(so[-1]):.4f}, fp32 end {float(st[-1]):.4f})"


@pytest.mark.skipif(not cuda, reason="needs a GPU")
def test_learners_train_as_one_cohort_and_evaluate():
    from myfyp_b200.learning.dataset import P2PFLDataset, RandomIIDPartitionStrategy
    from myfyp_b200.learning.torch_learner import TorchLearner
    from myfyp_b200.models import cnn_model
    from myfyp_b200.models.grouped_cnn import CohortTrainer

    CohortTrainer.reset()
    dev = torch.device("cuda", 0)
    parts = P2PFLDataset.synthetic_cifar10(5 * 500, 5 * 128, seed=1).generate_partitions(5, RandomIIDPartitionStrategy)
    learners = [TorchLearner(cnn_model(device=dev, seed=i, precision="tf32"), parts[i], f"n{i}", batch_size=64) for i in range(5)]
    for l in learners:
        l.set_epochs(4)
    out = {}

    def work(i):
        out[i, "before"] = learners[i].evaluate()
        learners[i].fit()
        out[i, "after"] = learners[i].evaluate()

    ths = [threading.Thread(target=work, args=(i,)) for i in range(5)]
    [t.start() for t in ths]
    [t.join(120) for t in ths]
    tr = next(iter(CohortTrainer._instances.values()))
    assert tr.batched == 15 and tr.batches < 15, f"requests were not batched: {tr.batches} launches for {tr.batched} requests"
    for i in range(5):
        assert out[i, "after"]["test_loss"] < out[i, "before"]["test_loss"], out
        assert out[i, "after"]["test_metric"] > out[i, "before"]["test_metric"] + 0.15, out
    # the native evaluation agrees with the module's own forward
    m = learners[0].get_model().get_model()
    x, y = parts[0].export(__import__("myfyp_b200.learning.dataset.p2pfl_dataset", fromlist=["TensorExportStrategy"]).TensorExportStrategy, train=False, device=dev)
    with torch.no_grad():
        logits = m(x)
    assert abs(float(F.cross_entropy(logits, y)) - out[0, "after"]["test_loss"]) < 5e-3
    assert abs(float((logits.argmax(1) == y).float().mean()) - out[0, "after"]["test_metric"]) < 0.02
    CohortTrainer.reset()
