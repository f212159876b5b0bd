#!/usr/bin/env python
"""ncu target for the multi-GPU reductions: ONE process, two GPUs, the K-way pull kernels reading half of their inputs from the
peer GPU over NVLink.  None of these kernels waits for another launch, so ncu's kernel replay is safe (the flag-synchronised
two-shot / NVLS collectives are NOT captured this way).

    python tools/ncu_peer_target.py                       # plain run first
    ncu --set full --clock-control none --import-source on -k "regex:fedavg|median" -s 4 -c 4 \
        --metrics <peer / NVLink counters found by `ncu --query-metrics`> -o gpurun_out/peer_pull python tools/ncu_peer_target.py

4 warm-up launches, then: fedavg (local only), fedavg (2 local + 2 peer), median (2 + 2), median sharded (its GPU-0 half)."""

from __future__ import annotations

import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch  # noqa: E402

from myfyp_b200 import ops  # noqa: E402

n = int(os.environ.get("PEER_N", str(16 << 20)))  # 64 MB per model
d0, d1 = torch.device("cuda", 0), torch.device("cuda", 1)
local = [torch.randn(n, device=d0) for _ in range(4)]
remote = [torch.randn(n, device=d1) for _ in range(2)]
out = torch.empty(n, device=d0)
w4 = [0.25] * 4
for _ in range(2):  # warm-up: 4 launches
    ops.fedavg(local, w4, device=d0, out=out)
    ops.fedavg(local[:2] + remote, w4, device=d0, out=out)
torch.cuda.synchronize()


def timed(fn, name: str, bytes_moved: int) -> None:
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    fn()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    print(f"{name:44s} {1e3 * ms:9.1f} us   {bytes_moved / ms / 1e6:8.1f} GB/s of inputs", flush=True)


timed(lambda: ops.fedavg(local, w4, device=d0, out=out), "fedavg 4 local", 4 * 4 * n)
timed(lambda: ops.fedavg(local[:2] + remote, w4, device=d0, out=out), "fedavg 2 local + 2 peer", 4 * 4 * n)
timed(lambda: ops.median(local[:2] + remote, device=d0, out=out, shard=False), "median 2 local + 2 peer, one launch", 4 * 4 * n)
timed(lambda: ops.median(local[:2] + remote, device=d0, out=out, shard=True), "median 2 local + 2 peer, sharded over 2 GPUs", 4 * 4 * n)
ref = (local[0] + local[1] + remote[0].to(d0) + remote[1].to(d0)) * 0.25
ops.fedavg(local[:2] + remote, w4, device=d0, out=out)
torch.cuda.synchronize()
print("max err", float((out - ref).abs().max()))
