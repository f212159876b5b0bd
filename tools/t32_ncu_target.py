#!/usr/bin/env python
"""Smallest program that runs the node-batched TF32 step: 1 refresh + 3 training steps of G=8 slots (ncu target).

    ncu --set full --import-source on -k regex:t32 -s 27 -c 13 -o gpurun_out/t32_step python tools/t32_ncu_target.py
(launch 0 = refresh, 1..26 = two warm-up steps, 27..39 = the step that is captured)"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from myfyp_b200.models import cnn_model  # noqa: E402
from myfyp_b200.models import grouped_cnn as gc  # noqa: E402

G = int(os.environ.get("T32_G", "8"))
dev = torch.device("cuda", 0)
models = [cnn_model(device=dev, seed=g, precision="tf32") for g in range(G)]
plan = gc.CnnPlan(models[0].get_model().grouped_spec, models[0].layout, dev, G, 64)
n = 1024
slots = []
for m in models:
    nt = m.layout.n_trainable
    slots.append(dict(slab=m.slab, m=torch.zeros(nt, device=dev), v=torch.zeros(nt, device=dev), adam_state=torch.zeros(4, device=dev), x=torch.randn(n, 3, 32, 32, device=dev),
                      y=torch.randint(0, 10, (n,), device=dev), perm=torch.randperm(n, device=dev), losses=torch.zeros(16, device=dev), n_samples=n, lr=1e-3))
plan.bind(slots)
plan.refresh()
for s in range(3):
    plan.train_step(s)
torch.cuda.synchronize()
print("ok", [float(s["losses"][2]) for s in slots][:2])
