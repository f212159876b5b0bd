"""FedMedian on the GPU with ResNet-18 nodes: does the federation learn without / with sign-flip attackers? (per-round accuracies)"""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from benchmarks import configs
from myfyp_b200.learning.aggregators import FedAvg, FedMedian
from myfyp_b200.management.logger import logger
from myfyp_b200.settings import Settings
import builtins
for name, agg, byz in (("median-clean", FedMedian, 0), ("median-attack", FedMedian, 2), ("fedavg-clean", FedAvg, 0)):
    out = configs.run(name, 16, "resnet18", agg, "random", 8, 256, 8, "cuda", byzantine=byz)
    glob = logger.get_global_logs().get(Settings.EXPERIMENT_NAME, {})
    accs = {k.split("-")[-1]: [round(v, 2) for _, v in m["test_metric"]] for k, m in glob.items() if "test_metric" in m}
    print(name, json.dumps(accs), flush=True)
