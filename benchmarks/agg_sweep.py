#!/usr/bin/env python
"""BASELINE.json config 5: aggregation bandwidth sweep, model size 1 MB - 1 GB at G = 2/4/8 GPUs.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node G --master-addr 127.0.0.1 benchmarks/agg_sweep.py [--sizes-mb 1 16 64 256 1024]

Every rank holds one fp32 model of the given size; after the operation every rank holds the sample-weighted average.
Arms:
  oneshot   our A1 kernel: every GPU pulls all G slabs (G-1 of them over NVLink through IPC mappings) and reduces in one pass
  twoshot   our A2 kernel: fused reduce-scatter + all-gather over peer memory with in-kernel flags
  nccl      BASELINE: pre-weight + torch.distributed NCCL all_reduce
  numpy     BASELINE: the reference's FedAvg math (fedavg.py:61-69) on G NumPy models on the host (rank 0, sizes <= 256 MB)
Device time (CUDA events, max over ranks), median of the iterations, inputs far larger than L2 for >= 64 MB (smaller
sizes are latency numbers; L2 is flushed before each iteration anyway).
Roofline: bytes that must cross NVLink into one GPU / 770 GB/s measured peer-copy bandwidth (B200_PROFILING.md):
oneshot (G-1)*S, twoshot 2*(G-1)/G*S.
"""

from __future__ import annotations

import argparse
import json
import os
import statistics
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

NVLINK_GBS = 770.0


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--sizes-mb", type=float, nargs="*", default=[1, 4, 16, 64, 256, 1024])
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--out", default=None)
    ap.add_argument("--no-nvls", action="store_true")
    a = ap.parse_args()
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("cpu:gloo,cuda:nccl", device_id=dev)
    from myfyp_b200 import ops
    from myfyp_b200.parallel.collective import NvlsFedAvg, TwoShotFedAvg, nvls_available

    use_nvls = world > 1 and nvls_available() and not a.no_nvls

    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    samples = [float(10 + 3 * r) for r in range(world)]
    weights = [s / sum(samples) for s in samples]
    rows = []

    def timed(fn, iters):
        ts = []
        for it in range(iters + 3):
            flush.zero_()
            torch.cuda.synchronize(dev)
            dist.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            torch.cuda.synchronize(dev)
            t = torch.tensor([e0.elapsed_time(e1)], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            if it >= 3:
                ts.append(float(t.item()))
        return statistics.median(ts)

    for mb in a.sizes_mb:
        n = int(mb * (1 << 20) / 4) // 64 * 64
        size = 4 * n
        coll = TwoShotFedAvg(n, rank, world, dev)
        torch.manual_seed(rank)
        coll.src.copy_(torch.randn(n, device=dev))
        x = coll.src.clone()
        out1 = torch.empty(n, device=dev)
        dist.barrier()
        # correctness of all three against each other
        ops.fedavg(coll.srcs, weights, out=out1)
        torch.cuda.synchronize(dev)
        dist.barrier()
        o2 = coll(weights).clone()
        torch.cuda.synchronize(dev)
        y = x * weights[rank]
        dist.all_reduce(y)
        err12 = float((out1 - o2).abs().max())
        err1n = float((out1 - y).abs().max())
        iters = a.iters if mb <= 256 else max(3, a.iters // 2)
        t_one = timed(lambda: ops.fedavg(coll.srcs, weights, out=out1), iters)
        t_two = timed(lambda: coll(weights), iters)

        def nccl():
            yy = x * weights[rank]
            dist.all_reduce(yy)

        t_nccl = timed(nccl, iters)
        t_nvls, err1v = None, None
        if use_nvls:
            # in-switch reduction: pre-weight into the bound buffer (the same elementwise pass the NCCL arm pays), one kernel
            nv = NvlsFedAvg(n, rank, world, dev)

            def nvls():
                torch.mul(x, weights[rank], out=nv.src)
                nv(max_ctas=296 if mb >= 64 else 0)  # the sweep owns the GPU: two resident CTAs per SM for the big sizes

            nvls()
            torch.cuda.synchronize(dev)
            if nv.timed_out():
                raise RuntimeError("NVLS all-reduce timed out")
            err1v = float((out1 - nv.out).abs().max())
            t_nvls = timed(nvls, iters)
            nv.close()
            del nv
        row = {
            "gpus": world, "model_MB": mb,
            "oneshot_ms": round(t_one, 4), "twoshot_ms": round(t_two, 4), "nccl_ms": round(t_nccl, 4),
            "oneshot_algbw_GBs": round(size / t_one / 1e6, 1), "twoshot_algbw_GBs": round(size / t_two / 1e6, 1), "nccl_algbw_GBs": round(size / t_nccl / 1e6, 1),
            "oneshot_input_GBs": round(world * size / t_one / 1e6, 1), "twoshot_input_GBs": round(world * size / t_two / 1e6, 1),
            "oneshot_roofline_frac": round(((world - 1) * size / (NVLINK_GBS * 1e6)) / t_one, 3),
            "twoshot_roofline_frac": round((2 * (world - 1) / world * size / (NVLINK_GBS * 1e6)) / t_two, 3),
            "speedup_vs_nccl": round(t_nccl / min(t for t in (t_one, t_two, t_nvls) if t), 2), "max_abs_diff_one_vs_two": err12, "max_abs_diff_one_vs_nccl": err1n,
        }
        if t_nvls:
            # NVLink bytes per GPU with the reduction in the switch: S/G read-reduced + S/G multicast-stored on the sending side,
            # and (G-1)/G * S arriving - the receive direction is the bound
            row.update(nvls_ms=round(t_nvls, 4), nvls_algbw_GBs=round(size / t_nvls / 1e6, 1), nvls_roofline_frac=round(((world - 1) / world * size / (NVLINK_GBS * 1e6)) / t_nvls, 3),
                       nvls_vs_nccl=round(t_nccl / t_nvls, 2), max_abs_diff_one_vs_nvls=err1v)
        if rank == 0 and mb <= 256:
            import numpy as np

            models = [np.random.default_rng(g).standard_normal(n).astype(np.float32) for g in range(world)]
            t0 = time.perf_counter()
            acc = np.zeros_like(models[0])
            for m, s in zip(models, samples):
                acc = np.add(acc, m * s)  # the reference's loop, fedavg.py:64-66
            acc = np.divide(acc, sum(samples))
            t_np = (time.perf_counter() - t0) * 1e3
            row.update(numpy_ms=round(t_np, 2), numpy_input_GBs=round(world * size / t_np / 1e6, 2), speedup_vs_numpy=round(t_np / min(t_one, t_two), 1))
        rows.append(row)
        if rank == 0:
            print(json.dumps(row), flush=True)
        coll.close()
        del coll, x, out1, o2, y
        torch.cuda.empty_cache()
        dist.barrier()
    if rank == 0 and a.out:
        os.makedirs(os.path.dirname(os.path.abspath(a.out)), exist_ok=True)
        with open(a.out, "w") as f:
            json.dump(rows, f, indent=1)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
