#!/usr/bin/env python
"""Aggregation-step baselines for BASELINE.json configs 3 and 4 (ResNet-18-sized models, K = 8 contributors - the train-set size).

What the reference offers for these aggregators, on the same parameter values our kernels see:
  scaffold   the reference's own class, unmodified from baseline/_ref (p2pfl/learning/aggregators/scaffold.py:70-122), on NumPy lists
  fedmedian  the reference's class raises NotImplementedError (fedmedian.py:47); baseline = its intended math (:55-65),
             np.median(np.array(layer_weights), axis=0) per layer
  krum       absent from the reference; baseline = a NumPy restatement (pairwise squared distances per layer pair, scores, argmin)
  fedavg     the reference's FedAvg (fedavg.py:41-77) for scale
against our device kernels (b2_scaffold_server, b2_median, b2_krum_gram + b2_krum_select + b2_fedavg, b2_fedavg) on slabs in HBM.
Times are per aggregation (median of `--iters`); ours are CUDA-event times.  One JSON line per aggregator.

    python benchmarks/agg_baselines.py [--k 8] [--model resnet18] [--out benchmarks/results/agg_baselines.json]
"""

from __future__ import annotations

import argparse
import json
import os
import statistics
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
REF = os.path.join(ROOT, "baseline", "_ref")


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--k", type=int, default=8)
    ap.add_argument("--model", default="resnet18", choices=["resnet18", "cnn", "mlp"])
    ap.add_argument("--iters", type=int, default=5)
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    import numpy as np
    import torch

    from myfyp_b200 import ops
    from myfyp_b200.learning.aggregators import FedAvg, FedMedian, Krum, Scaffold
    from myfyp_b200.models import cnn_model, mlp_model, resnet18_model

    dev = torch.device("cuda", 0) if torch.cuda.is_available() else torch.device("cpu")
    mk = {"resnet18": resnet18_model, "cnn": cnn_model, "mlp": mlp_model}[a.model]
    base = mk(device=dev, seed=0)
    rng = np.random.default_rng(0)
    ref_params = base.get_parameters()
    models, np_models = [], []
    for k in range(a.k):
        params = [p + 0.01 * rng.standard_normal(p.shape).astype(np.float32) if p.dtype == np.float32 else p for p in ref_params]
        m = base.build_copy(params=params, num_samples=100 + 10 * k, contributors=[f"n{k}"])
        m.add_info("scaffold", {"delta_y_i": [0.01 * (k + 1) * np.ones_like(p, dtype=np.float32) for p in ref_params], "delta_c_i": [0.001 * np.ones_like(p, dtype=np.float32) for p in ref_params]})
        models.append(m)
        np_models.append(params)
    nbytes = 4 * base.layout.total

    def dev_time(fn):
        ts = []
        for _ in range(a.iters + 2):
            if dev.type == "cuda":
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                fn()
                e1.record()
                torch.cuda.synchronize()
                ts.append(e0.elapsed_time(e1))
            else:
                t0 = time.perf_counter()
                fn()
                ts.append(1e3 * (time.perf_counter() - t0))
        return statistics.median(ts[2:])

    def host_time(fn):
        ts = []
        for _ in range(max(2, a.iters // 2)):
            t0 = time.perf_counter()
            fn()
            ts.append(1e3 * (time.perf_counter() - t0))
        return statistics.median(ts)

    rows = []
    slabs = [m.slab for m in models]
    # ---------------- FedAvg
    cwd = os.getcwd()
    os.chdir(tempfile.mkdtemp(prefix="aggbase-"))
    sys.path.insert(0, REF)
    ref_ok = os.path.isdir(os.path.join(REF, "p2pfl"))
    if ref_ok:
        from p2pfl.learning.aggregators.fedavg import FedAvg as RefFedAvg
        from p2pfl.learning.aggregators.scaffold import Scaffold as RefScaffold
        from p2pfl.learning.frameworks.p2pfl_model import P2PFLModel as RefModel

        class NpModel(RefModel):
            def __init__(self, params, num_samples=None, contributors=None, additional_info=None):
                self._p = params
                self.num_samples, self.contributors, self.additional_info = num_samples or 0, contributors or [], additional_info or {}

            def get_parameters(self):
                return self._p

            def build_copy(self, **kw):
                return NpModel(kw.get("params", self._p), kw.get("num_samples"), kw.get("contributors"))

        # the reference's Scaffold accumulates in arrays of the parameters' own dtype and fails on integer state entries
        # (BatchNorm's int64 batch counter: "Cannot cast ufunc 'add' output"); it is given float32 copies of those
        as_f32 = lambda arrs: [np.asarray(a, dtype=np.float32) for a in arrs]  # noqa: E731
        ref_models = [NpModel(as_f32(p), m.get_num_samples(), m.get_contributors(), {"scaffold": {k: as_f32(v) for k, v in m.get_info("scaffold").items()}}) for p, m in zip(np_models, models)]
        t_ref_avg = host_time(lambda: RefFedAvg().aggregate(ref_models))
        t_ref_scaf = host_time(lambda: RefScaffold().aggregate(ref_models))
    else:
        t_ref_avg = t_ref_scaf = None
    os.chdir(cwd)
    t_avg = dev_time(lambda: FedAvg().aggregate(models))
    rows.append({"aggregator": "fedavg", "ours_ms": t_avg, "reference_ms": t_ref_avg, "baseline": "reference FedAvg.aggregate (NumPy)"})
    # ---------------- Scaffold
    from myfyp_b200.parallel.slab import flat_from_wire

    for m in models:  # on the fabric the deltas are flat device slabs (what SCAFFOLDCallback produces), not per-layer host arrays
        m.add_info("scaffold", {k: flat_from_wire(v, m.layout, dev) for k, v in m.get_info("scaffold").items()})
    t_scaf = dev_time(lambda: Scaffold().aggregate(models))
    rows.append({"aggregator": "scaffold", "ours_ms": t_scaf, "reference_ms": t_ref_scaf, "baseline": "reference Scaffold.aggregate (NumPy), unmodified"})
    # ---------------- FedMedian
    def np_median():
        return [np.median(np.array([pm[i] for pm in np_models]), axis=0) for i in range(len(ref_params))]

    t_np_med = host_time(np_median)
    t_med = dev_time(lambda: FedMedian().aggregate(models))
    rows.append({"aggregator": "fedmedian", "ours_ms": t_med, "reference_ms": t_np_med, "baseline": "NumPy restatement of the reference's intended math (fedmedian.py:55-65)"})
    # ---------------- Krum
    def np_krum():
        flat = [np.concatenate([np.asarray(l, dtype=np.float32).ravel() for l in pm]) for pm in np_models]
        k = len(flat)
        d = np.zeros((k, k))
        for i in range(k):
            for j in range(i + 1, k):
                diff = flat[i] - flat[j]
                d[i, j] = d[j, i] = float(diff @ diff)
        scores = [np.sort(np.delete(d[i], i))[: max(1, k - 2 - 2)].sum() for i in range(k)]
        return flat[int(np.argmin(scores))]

    t_np_krum = host_time(np_krum)
    t_krum = dev_time(lambda: Krum(f=2, m=1).aggregate(models))
    rows.append({"aggregator": "krum", "ours_ms": t_krum, "reference_ms": t_np_krum, "baseline": "NumPy restatement (Krum is not in the reference)"})
    for r in rows:
        r.update(model=a.model, k=a.k, model_MB=round(nbytes / 1e6, 1), device=str(dev), speedup=(round(r["reference_ms"] / r["ours_ms"], 1) if r["reference_ms"] else None),
                 ours_input_GBs=round(a.k * nbytes / r["ours_ms"] / 1e6, 1))
        r["ours_ms"] = round(r["ours_ms"], 4)
        r["reference_ms"] = round(r["reference_ms"], 2) if r["reference_ms"] else None
        print(json.dumps(r), flush=True)
    if a.out:
        os.makedirs(os.path.dirname(os.path.abspath(a.out)), exist_ok=True)
        with open(a.out, "w") as f:
            json.dump(rows, f, indent=1)


if __name__ == "__main__":
    main()
