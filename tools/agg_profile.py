"""Launch every aggregation / optimizer / loss kernel once at a representative size (for one `ncu --set full` pass)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from myfyp_b200 import ops
from myfyp_b200.parallel.collective import TwoShotFedAvg

dev = torch.device("cuda:0")
n = 11 * (1 << 20)  # ~ ResNet-18 sized slab (44 MB)
k = 8
slabs = [torch.randn(n, device=dev) for _ in range(k)]
out = torch.empty(n, device=dev)
w = [1.0 / k] * k
coll = TwoShotFedAvg.emulated(n, 4, dev)
for r in range(4):
    coll.srcs[r].copy_(slabs[r])
m_, v_, st = torch.zeros(n, device=dev), torch.zeros(n, device=dev), torch.zeros(4, device=dev)
p16 = torch.zeros(n, dtype=torch.bfloat16, device=dev)
lg = torch.randn(4096, 10, device=dev); lab = torch.randint(0, 10, (4096,), device=dev); acc = torch.zeros(2, device=dev); conf = torch.zeros(10, 10, device=dev)
for it in range(2):  # first pass = warm-up (ncu is told to skip it)
    ops.fedavg(slabs, w, out=out)
    ops.accumulate(out, slabs[1], 0.5, beta=1.0)
    ops.pull_copy(out, slabs[2])
    ops.median(slabs, out=out)
    ops.median(slabs, trim=1, out=out)
    d = ops.pairwise_sqdist(slabs)
    ops.krum_select(d, 1, 3)
    ops.scaffold_server_update(slabs[0], slabs[1], slabs[2:5], slabs[5:8], [0.3, 0.3, 0.4], 1.0 / 3)
    ops.scaffold_finalize(slabs[0], slabs[1], slabs[2], slabs[3], slabs[4], 0.098)
    coll([0.25] * 4)
    ops.adam_step(slabs[0], slabs[1], m_, v_, 0, dev_state=st, step_offset=it + 1, p16=p16, prox_mu=0.01, prox_ref=slabs[2], cv_global=slabs[3], cv_local=slabs[4])
    ops.sgd_step(slabs[0], slabs[1], m_, momentum=0.9)
    ops.cast_slab(slabs[0], p16)
    ops.cross_entropy_eval(lg, lab, acc)
    ops.confusion(lg, lab, conf)
    torch.cuda.synchronize()
print("agg profile ok")
