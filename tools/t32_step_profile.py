#!/usr/bin/env python
"""Per-kernel device time of the node-batched TF32 CNN step (csrc/train_tf32.cu), CUDA events, warm, G slots active.

    python tools/t32_step_profile.py [--groups 8] [--batch 64] [--reps 30] [--json out.json]

Each kernel is launched alone `reps` times between two events (L2 stays warm, as inside a step); the whole step and a captured
98-step epoch graph are timed too.  The table is what profiles/t32_step.md is made from."""

from __future__ import annotations

import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch  # noqa: E402


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--groups", type=int, default=8)
    ap.add_argument("--active", type=int, default=0)
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--reps", type=int, default=30)
    ap.add_argument("--samples", type=int, default=6250)
    ap.add_argument("--json", default="")
    a = ap.parse_args()
    from myfyp_b200.models import cnn_model
    from myfyp_b200.models import grouped_cnn as gc

    dev = torch.device("cuda", 0)
    G, act = a.groups, (a.active or a.groups)
    models = [cnn_model(device=dev, seed=g, precision="tf32") for g in range(act)]
    plan = gc.CnnPlan(models[0].get_model().grouped_spec, models[0].layout, dev, G, a.batch)
    steps = (a.samples + a.batch - 1) // a.batch
    slots = []
    for m in models:
        nt = m.layout.n_trainable
        slots.append(dict(slab=m.slab, m=torch.zeros(nt, device=dev), v=torch.zeros(nt, device=dev), adam_state=torch.zeros(4, device=dev),
                          x=torch.randn(a.samples, 3, 32, 32, device=dev), y=torch.randint(0, 10, (a.samples,), device=dev), perm=torch.randperm(a.samples, device=dev),
                          losses=torch.zeros(steps, device=dev), n_samples=a.samples, lr=1e-3))
    plan.bind(slots + [None] * (G - act))
    plan.refresh()
    for s in range(3):
        plan.train_step(s)
    torch.cuda.synchronize()
    seq = [("conv0_fwd", gc.RUN_CONV0_FWD), ("conv1_fwd", gc.RUN_CONV_FWD + 1), ("conv2_fwd", gc.RUN_CONV_FWD + 2), ("fc1_fwd", gc.RUN_FC1_FWD), ("head", gc.RUN_HEAD_TRAIN),
           ("fc1_wgrad*", gc.RUN_FC1_WGRAD), ("fc1_dgrad", gc.RUN_FC1_DGRAD), ("conv2_wgrad*", gc.RUN_CONV_WGRAD + 2), ("conv2_dgrad", gc.RUN_CONV_DGRAD + 2),
           ("conv1_wgrad*", gc.RUN_CONV_WGRAD + 1), ("conv1_dgrad", gc.RUN_CONV_DGRAD + 1), ("conv0_wgrad", gc.RUN_CONV0_WGRAD), ("adam", gc.RUN_ADAM)]
    rows = {}
    side = torch.cuda.Stream()
    for name, which in seq:
        plan.run(which, 1)
        torch.cuda.synchronize()
        # `reps` copies of the kernel in one graph: the replay has no host launch overhead between them
        gk = torch.cuda.CUDAGraph()
        with torch.cuda.stream(side):
            with torch.cuda.graph(gk, stream=side):
                for _ in range(a.reps):
                    plan.run(which, 1)
            gk.replay()
            side.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            gk.replay()
            e1.record()
            side.synchronize()
        rows[name] = 1e3 * e0.elapsed_time(e1) / a.reps
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for r in range(a.reps):
        plan.train_step(r % steps)
    e1.record()
    torch.cuda.synchronize()
    rows["step (eager, 13 launches)"] = 1e3 * e0.elapsed_time(e1) / a.reps
    g = torch.cuda.CUDAGraph()
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        with torch.cuda.graph(g, stream=st):
            for s in range(steps):
                plan.train_step(s)
        g.replay()
        st.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(3):
            g.replay()
        e1.record()
        st.synchronize()
    rows[f"step inside the {steps}-step epoch graph"] = 1e3 * e0.elapsed_time(e1) / 3 / steps
    # the same graph after the GPU sat idle (what a federation's round looks like: votes and gossip between two fits)
    import time

    scratch = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    with torch.cuda.stream(st):
        for idle_ms, warm in ((2, 0), (10, 0), (30, 0), (100, 0), (30, 1), (30, 5)):
            ts = []
            for _ in range(3):
                st.synchronize()
                time.sleep(idle_ms / 1e3)
                t0 = time.perf_counter()
                while warm and time.perf_counter() - t0 < warm / 1e3:  # keep the GPU busy for `warm` ms first
                    scratch.zero_()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                g.replay()
                e1.record()
                st.synchronize()
                ts.append(1e3 * e0.elapsed_time(e1) / steps)
            rows[f"step in the epoch graph after {idle_ms} ms idle" + (f" + {warm} ms of memsets" if warm else "")] = sorted(ts)[1]
    crit = sum(v for k, v in rows.items() if not k.endswith("*") and not k.startswith("step"))
    print(f"groups={G} active={act} batch={a.batch}   (us per launch, all active slots together; * = side branch)")
    for k, v in rows.items():
        print(f"  {k:42s} {v:9.1f}")
    print(f"  {'sum of the critical-path kernels':42s} {crit:9.1f}")
    if a.json:
        with open(a.json, "w") as f:
            json.dump({"groups": G, "active": act, "batch": a.batch, "us": rows, "critical_path_us": crit}, f, indent=1)


if __name__ == "__main__":
    main()
