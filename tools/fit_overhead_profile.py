#!/usr/bin/env python
"""Where does `TorchLearner.fit()` spend host time around the epoch graph?  One node, the bench's CNN and data shape, cProfile over
20 fits (after 3 warm-up fits), wall time per fit next to the device time of the graph.

    python tools/fit_overhead_profile.py [--samples 6250] [--fits 20]"""
from __future__ import annotations

import argparse
import cProfile
import os
import pstats
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch  # noqa: E402


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--samples", type=int, default=6250)
    ap.add_argument("--fits", type=int, default=20)
    a = ap.parse_args()
    from myfyp_b200.learning.dataset.p2pfl_dataset import P2PFLDataset
    from myfyp_b200.learning.torch_learner import TorchLearner
    from myfyp_b200.management.logger import logger
    from myfyp_b200.models import cnn_model

    dev = torch.device("cuda", 0)
    data = P2PFLDataset.synthetic((3, 32, 32), 10, a.samples, 1250, seed=0)
    learner = TorchLearner(cnn_model(device=dev, seed=0), data, "n0", batch_size=64)
    learner.set_epochs(1)
    for _ in range(3):
        learner.fit()
        learner.evaluate()
    torch.cuda.synchronize()
    pr = cProfile.Profile()
    t0 = time.perf_counter()
    pr.enable()
    for _ in range(a.fits):
        learner.fit()
    pr.disable()
    wall = 1e3 * (time.perf_counter() - t0) / a.fits
    dt = logger.get_device_times()
    graph = dt.get("cohort/fit_graph_on_device", [0.0])
    print(f"wall per fit {wall:.2f} ms; graph on device {sum(graph[-a.fits:]) / a.fits:.2f} ms; laps: " +
          ", ".join(f"{k.split('/')[1]} {sum(v[-a.fits:]) / a.fits:.2f}" for k, v in dt.items() if k.startswith("cohort/fit_")))
    t0 = time.perf_counter()
    for _ in range(a.fits):
        learner.evaluate()
    print(f"wall per evaluate {1e3 * (time.perf_counter() - t0) / a.fits:.2f} ms")
    pstats.Stats(pr).sort_stats("cumulative").print_stats(28)


if __name__ == "__main__":
    main()
