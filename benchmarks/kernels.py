#!/usr/bin/env python
"""Per-kernel microbenchmarks with roofline fractions (one GPU).

    python benchmarks/kernels.py [fedavg median krum gemm adam xent copy] [--iters 20] [--json out.json]

Timing: CUDA events on the launching stream, >= 3 warm-ups, L2 flushed (256 MiB write) before every timed iteration,
median of the iterations.  Fractions are of the MEASURED peaks in MEASURED_PEAKS.json (copy bandwidth / cuBLAS bf16),
fallback 6.65 TB/s and 1.59 PFLOP/s ("of fallback") when the file is absent.
"""

from __future__ import annotations

import argparse
import json
import os
import statistics
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

from myfyp_b200 import ops  # noqa: E402


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return p["hbm_gbs"], p["bf16_tflops"], "measured"
    except Exception:
        return 6650.0, 1590.0, "fallback"


class Timer:
    def __init__(self, dev):
        self.dev = dev
        self.flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def run(self, fn, iters=20, warmup=3):
        for _ in range(warmup):
            fn()
        torch.cuda.synchronize(self.dev)
        ts = []
        for _ in range(iters):
            self.flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            torch.cuda.synchronize(self.dev)
            ts.append(e0.elapsed_time(e1))
        return statistics.median(ts)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("which", nargs="*", default=["copy", "fedavg", "median", "krum", "gemm", "adam", "xent"])
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--json", default=None)
    ap.add_argument("--quick", action="store_true", help="one size per kernel (for ncu captures)")
    a = ap.parse_args()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    hbm, tfl, tag = peaks()
    T = Timer(dev)
    rows = []

    def report(name, cfg, ms, bytes_=None, flops=None):
        r = {"kernel": name, "config": cfg, "ms": round(ms, 4)}
        if bytes_ is not None:
            gbs = bytes_ / ms / 1e6
            r.update(gbs=round(gbs, 1), frac_hbm=round(gbs / hbm, 3))
        if flops is not None:
            t = flops / ms / 1e9
            r.update(tflops=round(t, 2), frac_bf16=round(t / tfl, 4))
        r["peaks"] = tag
        rows.append(r)
        print(json.dumps(r), flush=True)

    mb = 1 << 20
    if "copy" in a.which:
        for n in ([64 * mb] if a.quick else [mb // 4, 16 * mb, 64 * mb, 256 * mb]):
            src, dst = torch.randn(n, device=dev), torch.empty(n, device=dev)
            ms = T.run(lambda: ops.pull_copy(dst, src), a.iters)
            report("pull_copy (A4)", {"MB": 4 * n // mb}, ms, bytes_=8 * n)
    if "fedavg" in a.which:
        for k, n in ([(8, 64 * mb)] if a.quick else [(2, 235264), (8, 235264), (8, 4 * mb), (4, 64 * mb), (8, 64 * mb), (16, 16 * mb), (32, 11 * mb), (64, 4 * mb)]):
            slabs = [torch.randn(n, device=dev) for _ in range(k)]
            out = torch.empty(n, device=dev)
            w = [1.0 / k] * k
            ms = T.run(lambda: ops.fedavg(slabs, w, out=out), a.iters)
            report("fedavg K-way (A1)", {"K": k, "MB_per_model": round(4 * n / mb, 2)}, ms, bytes_=4 * n * (k + 1))
    if "median" in a.which:
        for k, n in ([(8, 16 * mb)] if a.quick else [(4, 16 * mb), (8, 16 * mb), (16, 11 * mb), (32, 11 * mb), (64, 4 * mb)]):
            slabs = [torch.randn(n, device=dev) for _ in range(k)]
            out = torch.empty(n, device=dev)
            ms = T.run(lambda: ops.median(slabs, out=out), a.iters)
            report("median K-way (A5)", {"K": k, "MB_per_model": round(4 * n / mb, 2)}, ms, bytes_=4 * n * (k + 1))
            ms = T.run(lambda: ops.median(slabs, trim=1 if k > 2 else 0, out=out), a.iters)
            report("trimmed mean (A5)", {"K": k, "MB_per_model": round(4 * n / mb, 2)}, ms, bytes_=4 * n * (k + 1))
    if "krum" in a.which:
        for k, n, p in ([(32, 11 * mb, 3)] if a.quick else [(8, 11 * mb, 3), (32, 11 * mb, 3), (64, 11 * mb, 3), (32, 11 * mb, 1), (8, 235264, 3)]):
            slabs = [torch.randn(n, device=dev) for _ in range(k)]
            ms = T.run(lambda: ops.pairwise_sqdist(slabs, passes=p), a.iters)
            kp = (k + 15) // 16 * 16
            report("krum gram tcgen05 (A6)", {"K": k, "MB_per_model": round(4 * n / mb, 2), "passes": p}, ms, bytes_=4 * n * (k + 1), flops=2.0 * 128 * kp * n * p)
    if "gemm" in a.which:
        for m, nn, kk in ([(4096, 4096, 4096)] if a.quick else [(64, 256, 784), (1024, 256, 784), (4096, 4096, 4096), (8192, 8192, 8192), (16384, 64, 288), (256, 784, 1024)]):
            A = torch.randn(m, kk, device=dev).to(torch.bfloat16)
            B = torch.randn(nn, kk, device=dev).to(torch.bfloat16)
            out = torch.empty(m, nn, dtype=torch.bfloat16, device=dev)
            ms = T.run(lambda: ops.gemm_bf16(A, B, m, nn, kk, out_bf16=out), a.iters)
            report("gemm bf16 tcgen05+TMA (T1)", {"M": m, "N": nn, "K": kk}, ms, bytes_=2 * (m * kk + nn * kk + m * nn), flops=2.0 * m * nn * kk)
    if "adam" in a.which:
        for n in ([64 * mb] if a.quick else [235264, 11 * mb, 64 * mb]):
            p, g, m_, v = (torch.randn(n, device=dev) for _ in range(4))
            v.abs_()
            ms = T.run(lambda: ops.adam_step(p, g, m_, v, 3), a.iters)
            report("fused adam (T6)", {"M_params": round(n / 1e6, 2)}, ms, bytes_=4 * n * 8)
    if "xent" in a.which:
        for b, c in [(64, 10), (4096, 10), (4096, 1000)]:
            lg = torch.randn(b, c, device=dev)
            lab = torch.randint(0, c, (b,), device=dev)
            acc = torch.zeros(2, device=dev)
            ms = T.run(lambda: ops.cross_entropy_eval(lg, lab, acc), a.iters)
            report("softmax-xent (T3/T8)", {"B": b, "C": c}, ms, bytes_=4 * b * c)
    if a.json:
        with open(a.json, "w") as f:
            json.dump(rows, f, indent=1)


if __name__ == "__main__":
    main()
