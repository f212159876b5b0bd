"""Scaffold vs plain FedAvg with the same SGD clients on ResNet-18 nodes (config 4's model): is a low accuracy the optimizer's pace
or the control variates?  One JSON line per variant."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from benchmarks.configs import run
from myfyp_b200.learning.aggregators import FedAvg, Scaffold
from myfyp_b200.management.logger import logger
logger.set_level("ERROR")
n = int(sys.argv[1]) if len(sys.argv) > 1 else 32
rounds = int(sys.argv[2]) if len(sys.argv) > 2 else 12
for nm, agg, kw in (("fedavg-sgd0.1", FedAvg, {"optimizer": "sgd", "lr": 0.1}), ("scaffold-sgd0.1", lambda: Scaffold(global_lr=1.0, variate_scaling="paper"), {"optimizer": "sgd", "lr": 0.1}),
                    ("scaffold-sgd0.03", lambda: Scaffold(global_lr=1.0, variate_scaling="paper"), {"optimizer": "sgd", "lr": 0.03})):
    out = run(nm, n, "resnet18", agg, "random", rounds, 256, 8, "cuda", learner_kw=kw)
    print("RESULT", json.dumps({k: out[k] for k in ("scenario", "nodes", "gpus", "rounds", "honest_acc_mean", "honest_acc_min", "rounds_per_s")}), flush=True)
