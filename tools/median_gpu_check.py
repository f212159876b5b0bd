"""FedMedian on the GPU with ResNet-18 nodes under sign-flip attackers: per-round accuracies + final evaluation."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from benchmarks import configs
from myfyp_b200.learning.aggregators import FedMedian, TrimmedMean
from myfyp_b200.management.logger import logger
from myfyp_b200.settings import Settings
n, byz = int(sys.argv[1]) if len(sys.argv) > 1 else 32, int(sys.argv[2]) if len(sys.argv) > 2 else 4
for name, agg in (("median-attack", FedMedian), ("trimmed2-attack", lambda: TrimmedMean(trim=2))):
    out = configs.run(name, n, "resnet18", agg, "random", 12, 256, 8, "cuda", byzantine=byz)
    glob = logger.get_global_logs().get(Settings.EXPERIMENT_NAME, {})
    accs = {k.split("-")[-1]: [round(v, 2) for _, v in m["test_metric"]] for k, m in glob.items() if "test_metric" in m}
    print(name, "final", out["honest_acc_mean"], out["honest_acc_min"], json.dumps(dict(list(accs.items())[:10])), flush=True)
