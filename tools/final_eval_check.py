"""After a short ResNet-18 federation on the GPU: logged per-round accuracy vs learner.evaluate() vs a plain torch eval-mode forward."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from myfyp_b200.communication.protocols.memory.memory_communication_protocol import InMemoryCommunicationProtocol
from myfyp_b200.learning.aggregators import FedAvg
from myfyp_b200.learning.dataset import P2PFLDataset, RandomIIDPartitionStrategy, TensorExportStrategy
from myfyp_b200.management.logger import logger
from myfyp_b200.models import resnet18_model
from myfyp_b200.node import Node
from myfyp_b200.settings import Settings
from myfyp_b200.utils.utils import set_fast_settings
from myfyp_b200.utils.topologies import TopologyFactory
set_fast_settings(train_set_size=4)
n = 4
data = P2PFLDataset.synthetic((3, 32, 32), 10, 256 * n, 64 * n, seed=0)
parts = data.generate_partitions(n, RandomIIDPartitionStrategy, seed=0)
nodes = [Node(resnet18_model(device="cuda:0", seed=i), parts[i], address=f"fe-{i}", aggregator=FedAvg(), protocol=InMemoryCommunicationProtocol, batch_size=64) for i in range(n)]
for x in nodes: x.start()
TopologyFactory.connect_nodes(TopologyFactory.generate_matrix("full", n), nodes)
time.sleep(0.5)
nodes[0].set_start_learning(rounds=8, epochs=1)
while not all(x.learning_workflow.finished for x in nodes): time.sleep(0.05)
glob = logger.get_global_logs().get(Settings.EXPERIMENT_NAME, {})
print("logged", {k[-1]: [round(v, 2) for _, v in m["test_metric"]] for k, m in glob.items() if "test_metric" in m})
for nd in nodes:
    r1 = nd.learner.evaluate()
    model = nd.learner.get_model(); module = model.get_model()
    x, y = nd.learner.data.export(TensorExportStrategy, train=False, device="cuda:0")
    module.eval()
    with torch.no_grad():
        acc_t = (module(x.float()).argmax(1) == y).float().mean().item()
    module.train()
    with torch.no_grad():
        acc_train_mode = (module(x.float()).argmax(1) == y).float().mean().item()
    bn = module.bn1
    print(nd.addr, "evaluate()", round(r1["test_metric"], 3), "torch eval-mode", round(acc_t, 3), "torch train-mode BN", round(acc_train_mode, 3),
          "bn1 running_var mean", round(bn.running_var.mean().item(), 4), "batches", int(bn.num_batches_tracked.item()))
os._exit(0)
