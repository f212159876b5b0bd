#!/usr/bin/env python
"""Out-of-bounds self-check for the hand-written kernels - the stand-in for compute-sanitizer, which is closed on the B200 pool
(gpurun answers: "compute-sanitizer is closed on this pool ... find a bad access with bounds checks and asserts of your own,
small cases, and a comparison with the CPU reference"; profiles/sanitizer/README.md).

Every tensor a kernel reads or writes is carved out of a larger allocation with GUARD floats on both sides:

* guards of INPUTS hold NaN - a read past either end poisons the result, which is compared with a PyTorch reference;
* guards of OUTPUTS (and of in-place state) hold a canary bit pattern that is compared bit for bit after the kernel - a stray
  store shows up as a changed canary.

Shapes are odd on purpose (tails of the vector loops, partial last batch).  Each family prints one "ok" line; any failure raises.

    python tools/bounds_check.py            # also run by tests/test_bounds_gpu.py
"""

from __future__ import annotations

import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch  # noqa: E402

GUARD = 1024  # floats on each side (4 KB: keeps the carved view 16-byte aligned)
CANARY = 0x7FA5C3E1  # a NaN payload nobody computes


class Guarded:
    """A float32 / int64 / uint8 view of ``n`` elements with guard bands; ``check()`` asserts the bands are untouched."""

    def __init__(self, n: int, dev, dtype=torch.float32, poison: bool = True) -> None:
        self.n, self.dtype = n, dtype
        if dtype == torch.float32:
            raw = torch.full((n + 2 * GUARD,), CANARY, dtype=torch.int32, device=dev)
            self.raw = raw
            self.t = raw[GUARD : GUARD + n].view(torch.float32)
        else:
            fill = -(1 << 40) if dtype == torch.int64 else 255  # an index far outside any table / an impossible label
            self.raw = torch.full((n + 2 * GUARD,), fill, dtype=dtype, device=dev)
            self.t = self.raw[GUARD : GUARD + n]
        self._fill = self.raw[:GUARD].clone(), self.raw[GUARD + n :].clone()

    def set(self, src: torch.Tensor) -> "Guarded":
        self.t.copy_(src.reshape(-1).to(self.t.device))
        return self

    def check(self, what: str) -> None:
        torch.cuda.synchronize()
        assert torch.equal(self.raw[:GUARD], self._fill[0]), f"{what}: store BEFORE the buffer"
        assert torch.equal(self.raw[GUARD + self.n :], self._fill[1]), f"{what}: store PAST the buffer"


def run(dev=None, verbose: bool = True) -> None:
    from myfyp_b200 import ops

    dev = dev or torch.device("cuda", 0)
    g = torch.Generator(device="cpu").manual_seed(0)
    say = print if verbose else (lambda *a, **k: None)

    def G(n, dtype=torch.float32):
        return Guarded(n, dev, dtype)

    def rnd(n):
        return G(n).set(torch.randn(n, generator=g))

    # ---- K-way reductions (csrc/aggregate.cu): odd length = scalar tails after the 16-byte loops
    n = 100_003
    ins = [rnd(n) for _ in range(5)]
    w = [0.1, 0.2, 0.3, 0.2, 0.2]
    out = G(n)
    ops.fedavg([i.t for i in ins], w, device=dev, out=out.t)
    ref = sum(wi * i.t.double() for wi, i in zip(w, ins)).float()
    torch.testing.assert_close(out.t, ref, rtol=1e-5, atol=1e-5)
    acc = G(n).set(torch.zeros(n))
    ops.accumulate(acc.t, ins[0].t, 0.5)
    torch.testing.assert_close(acc.t, 0.5 * ins[0].t)
    for b in ins + [out, acc]:
        b.check("fedavg / accumulate")
    say("ok fedavg / accumulate", flush=True)
    for k in (3, 4, 5, 9, 16, 17, 33, 64):
        many = [rnd(n) for _ in range(k)]
        out = G(n)
        ops.median([m.t for m in many], device=dev, out=out.t)
        st = torch.stack([m.t for m in many]).sort(dim=0).values
        torch.testing.assert_close(out.t, (st[(k - 1) // 2] + st[k // 2]) * 0.5, rtol=1e-6, atol=1e-6)
        tr = 1 if k < 5 else 2
        ops.median([m.t for m in many], trim=tr, device=dev, out=out.t)
        torch.testing.assert_close(out.t, st[tr : k - tr].mean(dim=0), rtol=1e-5, atol=1e-6)
        for b in many + [out]:
            b.check(f"median K={k}")
    say("ok median / trimmed mean (K = 3 ... 64)", flush=True)
    n4 = 100_004  # the Scaffold kernels take whole 16-byte groups (the slab layout guarantees it)
    x, c = rnd(n4), G(n4).set(torch.zeros(n4))
    x0 = x.t.clone()
    dys, dcs = [rnd(n4) for _ in range(3)], [rnd(n4) for _ in range(3)]
    ops.scaffold_server_update(x.t, c.t, [d.t for d in dys], [d.t for d in dcs], [1 / 3] * 3, 1 / 3)
    torch.testing.assert_close(x.t, x0 + sum(d.t for d in dys) / 3, rtol=1e-5, atol=1e-5)
    torch.testing.assert_close(c.t, sum(d.t for d in dcs) / 3, rtol=1e-5, atol=1e-5)
    c_i, dy, dc, y = G(n4).set(torch.zeros(n4)), G(n4), G(n4), rnd(n4)
    ops.scaffold_finalize(c_i.t, dy.t, dc.t, x.t, y.t, 0.1)
    torch.testing.assert_close(dy.t, y.t - x.t)
    torch.testing.assert_close(dc.t, (x.t - y.t) / 0.1, rtol=1e-5, atol=1e-5)
    for b in [x, c, c_i, dy, dc, y] + dys + dcs:
        b.check("scaffold")
    say("ok scaffold server / finalize", flush=True)
    # ---- Krum: tcgen05 Gram over guarded slabs + selection
    ins = [rnd(n4) for _ in range(5)]  # the Gram kernel takes whole 16-byte groups as well
    dist = ops.pairwise_sqdist([i.t for i in ins], device=dev)
    big = torch.stack([i.t.double() for i in ins])
    torch.testing.assert_close(dist.double(), torch.cdist(big, big) ** 2, rtol=1e-4, atol=1e-2)
    scores, order = ops.krum_select(dist, f=1, m=2)
    assert all(s == s for s in scores) and len(order) == 2
    for b in ins:
        b.check("krum gram")
    say("ok krum gram / select", flush=True)
    # ---- loss / metrics (csrc/xent.cu)
    B, C = 77, 10
    logits, labels = rnd(B * C), G(B, torch.int64).set(torch.randint(0, C, (B,), generator=g))
    accv, conf = G(2).set(torch.zeros(2)), G(C * C).set(torch.zeros(C * C))
    ops.cross_entropy_eval(logits.t.view(B, C), labels.t, accv.t)
    ops.confusion(logits.t.view(B, C), labels.t, conf.t.view(C, C))
    want = torch.nn.functional.cross_entropy(logits.t.view(B, C), labels.t, reduction="sum")
    torch.testing.assert_close(accv.t[0], want, rtol=1e-4, atol=1e-3)
    assert float(conf.t.sum()) == B
    for b in (logits, labels, accv, conf):
        b.check("xent / confusion")
    say("ok xent / confusion", flush=True)
    # ---- node-batched TF32 step (csrc/train_tf32.cu): 2 slots, 3 steps (the third is a partial batch), then evaluation.
    # Data set, labels, permutation, losses, slab and optimizer state all sit between guards; a read past the data set would
    # pull NaN pixels / an impossible label into the loss.
    from myfyp_b200.models import cnn_model
    from myfyp_b200.models import grouped_cnn as gc

    Gs, Bt, ns = 2, 64, 160
    models = [cnn_model(device=dev, seed=s, precision="tf32") for s in range(Gs)]
    plan = gc.CnnPlan(models[0].get_model().grouped_spec, models[0].layout, dev, Gs, Bt)
    slots, guards = [], []
    for m in models:
        nt, tot = m.layout.n_trainable, m.slab.numel()
        slab = G(tot).set(m.slab)
        mm, vv, st = G(nt).set(torch.zeros(nt)), G(nt).set(torch.zeros(nt)), G(4).set(torch.zeros(4))
        xs = rnd(ns * 3 * 32 * 32)
        ys = G(ns, torch.int64).set(torch.randint(0, 10, (ns,), generator=g))
        pm = G(ns, torch.int64).set(torch.randperm(ns, generator=g))
        ls = G(4).set(torch.zeros(4))
        guards += [slab, mm, vv, st, xs, ys, pm, ls]
        slots.append(dict(slab=slab.t, m=mm.t, v=vv.t, adam_state=st.t, x=xs.t.view(ns, 3, 32, 32), y=ys.t, perm=pm.t, losses=ls.t, n_samples=ns, lr=1e-3))
    plan.bind(slots)
    plan.refresh()
    for s in range(3):
        plan.train_step(s)
    torch.cuda.synchronize()
    for s in slots:
        assert bool(torch.isfinite(s["losses"][:3]).all()) and float(s["losses"][:3].min()) > 0.5, s["losses"]
        assert bool(torch.isfinite(s["slab"]).all())
        s["losses"].zero_()
        s["perm"] = None
    plan.bind(slots)
    plan.refresh()
    for s in range(3):
        plan.eval_step(s)
    torch.cuda.synchronize()
    for s in slots:
        assert bool(torch.isfinite(s["losses"][:2]).all()), s["losses"]
        assert 0 <= float(s["losses"][1]) <= ns
    for b in guards:
        b.check("tf32 step")
    plan.close()
    say("ok tf32 step (train, partial batch, eval)", flush=True)


if __name__ == "__main__":
    run()
    print("bounds check passed")
