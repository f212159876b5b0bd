"""Run one small CPU federation many times and report any run in which no node changes stage for `--stall` seconds
(a rendez-vous waiting for a time-out), with every node's bookkeeping at that moment.

    python tools/stall_hunt.py --nodes 6 --rounds 3 --topology line --train-set 4 --aggregator fedavg --iters 50
"""
import argparse, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from myfyp_b200.communication.protocols.memory.memory_communication_protocol import InMemoryCommunicationProtocol
from myfyp_b200.learning.aggregators import FedAvg, FedMedian, Krum, Scaffold
from myfyp_b200.learning.dataset import P2PFLDataset, RandomIIDPartitionStrategy
from myfyp_b200.management.logger import logger
from myfyp_b200.models import mlp_model
from myfyp_b200.node import Node
from myfyp_b200.settings import Settings
from myfyp_b200.utils.topologies import TopologyFactory
from myfyp_b200.utils.utils import set_test_settings, wait_convergence

ap = argparse.ArgumentParser()
ap.add_argument("--nodes", type=int, default=6); ap.add_argument("--rounds", type=int, default=3)
ap.add_argument("--topology", default="line"); ap.add_argument("--train-set", type=int, default=4)
ap.add_argument("--aggregator", default="fedavg"); ap.add_argument("--iters", type=int, default=30)
ap.add_argument("--stall", type=float, default=6.0); ap.add_argument("--fast", action="store_true")
ap.add_argument("--protocol", default="memory", choices=["memory", "grpc", "grpc-wire"], help="grpc-wire: real RPCs also between nodes of this process")
ap.add_argument("--kill", action="store_true", help="stop the first non-initiator node that enters TrainStage of round 1")
a = ap.parse_args()
if a.fast:
    from myfyp_b200.utils.utils import set_fast_settings
    set_fast_settings(a.train_set)
else:
    set_test_settings()
Settings.TRAIN_SET_SIZE = a.train_set
Settings.USE_SSL = False
if a.protocol == "grpc-wire":
    Settings.GRPC_LOCAL_SHORTCUT = False
from myfyp_b200.communication.protocols.grpc.grpc_communication_protocol import GrpcCommunicationProtocol
proto = InMemoryCommunicationProtocol if a.protocol == "memory" else GrpcCommunicationProtocol
Settings.TTL = max(Settings.TTL, a.nodes)
logger.set_level("ERROR")
agg = {"fedavg": FedAvg, "fedmedian": FedMedian, "krum": lambda: Krum(f=1), "scaffold": lambda: Scaffold(global_lr=1.0)}[a.aggregator]
n, stalls = a.nodes, 0
for it in range(a.iters):
    parts = P2PFLDataset.synthetic_mnist(96 * n, 32 * n, seed=it).generate_partitions(n, RandomIIDPartitionStrategy)
    nodes = [Node(mlp_model(seed=i), parts[i], protocol=proto, aggregator=agg(), batch_size=32) for i in range(n)]
    for nd in nodes:
        nd.start()
    TopologyFactory.connect_nodes(TopologyFactory.generate_matrix(a.topology, n, seed=it), nodes)
    wait_convergence(nodes, n - 1, only_direct=False, wait=20)
    killed = []
    if a.kill:
        import threading

        def watch():
            while not killed:
                for x in nodes[1:]:
                    if x.state.round == 1 and x.learning_workflow.history and x.learning_workflow.history[-1] == "TrainStage":
                        killed.append(x)
                        x.stop()
                        return
                time.sleep(0.002)

        threading.Thread(target=watch, daemon=True).start()
    t0 = time.time()
    nodes[0].set_start_learning(rounds=a.rounds, epochs=1)
    last, t_change = None, time.time()
    while not all(x.learning_workflow.finished for x in nodes if x not in killed):
        snap = [(x.state.round, x.learning_workflow.history[-1] if x.learning_workflow.history else None) for x in nodes if x not in killed]
        if snap != last:
            last, t_change = snap, time.time()
        if time.time() - t_change > a.stall:
            stalls += 1
            ix = {x.addr: i for i, x in enumerate(nodes)}
            nm = lambda v: ix.get(v, v)  # noqa: E731
            print(f"STALL in run {it} after {time.time() - t0:.1f} s; killed {[ix[k.addr] for k in killed]}")
            for i, x in enumerate(nodes):
                if x in killed:
                    continue
                st = x.state
                print(f"  {i}: round {st.round} {x.learning_workflow.history[-2:]} train_set {[nm(v) for v in st.train_set]} "
                      f"announced {{{', '.join(f'{nm(k)}: {[nm(c) for c in v]}' for k, v in st.models_aggregated.items())}}} "
                      f"nei_status { {nm(k): v for k, v in st.nei_status.items()} } ballots {[nm(k) for k in st.train_set_votes]} "
                      f"early {len(st.early_models)} held {[nm(c) for c in x.aggregator.get_aggregated_models()]} neighbours {len(x.get_neighbors(only_direct=False))}", flush=True)
            break
        time.sleep(0.02)
    for nd in nodes:
        try:
            nd.stop()
        except Exception:  # noqa: BLE001
            pass
    if stalls:
        break
print(f"{a.iters if not stalls else it + 1} runs, {stalls} stalled", flush=True)
os._exit(1 if stalls else 0)
