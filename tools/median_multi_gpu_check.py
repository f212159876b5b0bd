import json, sys
sys.path.insert(0, '.')
from benchmarks.configs import run
from myfyp_b200.learning.aggregators import FedMedian, FedAvg
from myfyp_b200.management.logger import logger
logger.set_level("ERROR")
import torch
for gpus in (1, 2):
    out = run(f"med{gpus}", 32, "resnet18", FedMedian, "random", 10, 256, 8, "cuda", byzantine=4, max_gpus=gpus)
    print("RESULT", json.dumps({k: out[k] for k in ("scenario", "gpus", "rounds_per_s", "honest_acc_mean", "honest_acc_min")}), flush=True)
