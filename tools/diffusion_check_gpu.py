import sys, time
import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from myfyp_b200.communication.protocols.memory.memory_communication_protocol import InMemoryCommunicationProtocol
from myfyp_b200.learning.aggregators import FedAvg
from myfyp_b200.learning.dataset import P2PFLDataset, RandomIIDPartitionStrategy
from myfyp_b200.management.logger import logger
from myfyp_b200.models import resnet18_model
from myfyp_b200.node import Node
from myfyp_b200.utils.utils import set_fast_settings
from myfyp_b200.utils.topologies import TopologyFactory
set_fast_settings(train_set_size=3)
from myfyp_b200.parallel.fabric import Fabric
if torch.cuda.device_count() > 1: Fabric.get().enable_peer_access()
n=8
data = P2PFLDataset.synthetic((3,32,32), 10, 256*n, 64*n, seed=0)
parts = data.generate_partitions(n, RandomIIDPartitionStrategy, seed=0)
nodes=[Node(resnet18_model(device=f"cuda:{i % torch.cuda.device_count()}", seed=i), parts[i], address=f"x-{i}", aggregator=FedAvg(), protocol=InMemoryCommunicationProtocol, batch_size=64) for i in range(n)]
for x in nodes: x.start()
TopologyFactory.connect_nodes(TopologyFactory.generate_matrix("random", n, p=0.3, seed=1), nodes)
dl=time.time()+20
while any(len(x.get_neighbors(only_direct=False))<n-1 for x in nodes) and time.time()<dl: time.sleep(0.05)
nodes[0].set_start_learning(rounds=8, epochs=1)
while not all(x.learning_workflow.finished for x in nodes): time.sleep(0.05)
ref=nodes[0].learner.get_model().slab
for x in nodes:
    s=x.learner.get_model().slab.to(ref.device)
    print(x.addr, "acc", round(x.learner.evaluate()["test_metric"],3), "max|diff to node0|", float((s-ref).abs().max()), "trainstage rounds", x.learning_workflow.history.count("TrainStage"))
import os; os._exit(0)
