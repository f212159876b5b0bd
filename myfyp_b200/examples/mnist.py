"""N nodes in one process learn the MNIST-shaped task (reference: p2pfl/examples/mnist.py).

Same switches as the reference example - ``--nodes --rounds --epochs --protocol --aggregator --topology
--measure_time --show_metrics`` - plus ``--model mlp|cnn|resnet18``, ``--byzantine K --attack ...`` (the fork author's
poisoning experiments, exp_SAVE3.txt) and ``--drop-node-at-round R``.  Data is synthetic (there is no network on the
target boxes); plots are written as JSON / CSV (no display), never shown.
"""

from __future__ import annotations

import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))


def parse() -> argparse.Namespace:
    ap = argparse.ArgumentParser(description=__doc__.split("\n\n")[0])
    ap.add_argument("--nodes", type=int, default=2)
    ap.add_argument("--rounds", type=int, default=2)
    ap.add_argument("--epochs", type=int, default=1)
    ap.add_argument("--protocol", default="memory", choices=["memory", "grpc", "unix"])
    ap.add_argument("--framework", default="pytorch", choices=["pytorch", "tensorflow", "flax"], help="tensorflow / flax: host-side bindings, need those frameworks installed (MLP only)")
    ap.add_argument("--aggregator", default="fedavg", choices=["fedavg", "fedmedian", "trimmedmean", "krum", "multikrum", "scaffold", "fedprox"])
    ap.add_argument("--topology", default="line", choices=["star", "full", "line", "ring", "random"])
    ap.add_argument("--model", default="mlp", choices=["mlp", "cnn", "resnet18"])
    ap.add_argument("--samples-per-node", type=int, default=600)
    ap.add_argument("--batch-size", type=int, default=32)
    ap.add_argument("--train-set-size", type=int, default=4)
    ap.add_argument("--partition", default="iid", choices=["iid", "dirichlet", "label-skew"])
    ap.add_argument("--byzantine", type=int, default=0, help="number of poisoning nodes (the last ones)")
    ap.add_argument("--attack", default="sign_flip", choices=["sign_flip", "gaussian", "scaled_update"])
    ap.add_argument("--drop-node-at-round", type=int, default=-1, help="stop the last node when it reaches this round")
    ap.add_argument("--seed", type=int, default=666)
    ap.add_argument("--measure_time", action="store_true")
    ap.add_argument("--show_metrics", action="store_true", help="print the metric stores and write them to --out")
    ap.add_argument("--out", default="mnist_metrics.json")
    ap.add_argument("--log-level", default="INFO")
    ap.add_argument("--profiling", action="store_true", help="cProfile the run and print the top functions (the reference uses yappi)")
    # flags of the reference's example (p2pfl/examples/mnist.py:80-87)
    ap.add_argument("--token", default="", help="API token of the P2PFL web services: stream logs and metrics there (needs --web-url)")
    ap.add_argument("--web-url", default="", help="base URL of the P2PFL web services")
    ap.add_argument("--reduced_dataset", action="store_true", help="tiny partitions (50 samples per node), just for testing")
    ap.add_argument("--use_scaffold", action="store_true", help="same as --aggregator scaffold")
    ap.add_argument("--disable_ray", action="store_true", help="reference flag: run the learners on per-node graphs instead of the per-GPU cohort trainer (the role Ray's actor pool plays there)")
    ap.add_argument("--plot", default=None, help="write the global metrics as a PNG (matplotlib Agg backend; never opens a window)")
    return ap.parse_args()


def _plot(global_logs, path: str) -> None:
    """Global metrics per node as a PNG - headless (Agg); skipped with a message when matplotlib is missing."""
    try:
        import matplotlib

        matplotlib.use("Agg")
        import matplotlib.pyplot as plt
    except ImportError:
        print("matplotlib is not installed: no plot written")
        return
    for exp, per_node in global_logs.items():
        names = sorted({k for m in per_node.values() for k in m})
        fig, axes = plt.subplots(1, max(len(names), 1), figsize=(5 * max(len(names), 1), 4), squeeze=False)
        for ax, name in zip(axes[0], names):
            for node, metrics in per_node.items():
                if name in metrics:
                    ax.plot([r for r, _ in metrics[name]], [v for _, v in metrics[name]], label=node)
            ax.set_title(f"{exp}: {name}")
            ax.set_xlabel("round")
        fig.savefig(path)
        plt.close(fig)
        print(f"plot written to {path}")
        return


def main() -> None:
    a = parse()
    from myfyp_b200.communication.protocols.grpc.grpc_communication_protocol import GrpcCommunicationProtocol
    from myfyp_b200.communication.protocols.memory.memory_communication_protocol import InMemoryCommunicationProtocol
    from myfyp_b200.learning.aggregators import FedAvg, FedMedian, FedProx, Krum, Scaffold, TrimmedMean
    from myfyp_b200.learning.dataset import DirichletPartitionStrategy, LabelSkewedPartitionStrategy, P2PFLDataset, RandomIIDPartitionStrategy
    from myfyp_b200.management.logger import logger
    from myfyp_b200.models import cnn_model, mlp_model, resnet18_model
    from myfyp_b200.node import Node
    from myfyp_b200.settings import Settings
    from myfyp_b200.utils.fault_injector import FaultInjector, make_byzantine
    from myfyp_b200.utils.seed import set_seed
    from myfyp_b200.utils.topologies import TopologyFactory
    from myfyp_b200.utils.utils import set_standalone_settings, wait_convergence, wait_to_finish

    set_standalone_settings()
    if a.use_scaffold:
        a.aggregator = "scaffold"
    if a.reduced_dataset:
        a.samples_per_node = 50
    if a.disable_ray:
        Settings.DISABLE_RAY = True  # the cohort trainer is this framework's actor pool (learning/simulation.py)
    if a.token and a.web_url:
        logger.connect_web(a.web_url, a.token)
    Settings.TRAIN_SET_SIZE = a.train_set_size
    Settings.USE_SSL = False
    logger.set_level(a.log_level)
    set_seed(a.seed)
    if a.topology in ("line", "ring", "random"):
        Settings.TTL = max(Settings.TTL, a.nodes)  # must exceed the graph diameter (reference: examples/mnist.py:153-156)

    shape = (28, 28) if a.model == "mlp" else (3, 32, 32)
    data = P2PFLDataset.synthetic(shape, 10, a.samples_per_node * a.nodes, max(100, a.samples_per_node // 5) * a.nodes, seed=a.seed)
    strat = {"iid": RandomIIDPartitionStrategy, "dirichlet": DirichletPartitionStrategy, "label-skew": LabelSkewedPartitionStrategy}[a.partition]
    parts = data.generate_partitions(a.nodes, strat, seed=a.seed)
    mk_model = {"mlp": mlp_model, "cnn": cnn_model, "resnet18": resnet18_model}[a.model]
    if a.framework == "tensorflow":
        from myfyp_b200.learning.frameworks.tensorflow.keras_model import MLP as KerasMLP
        from myfyp_b200.learning.frameworks.tensorflow.keras_model import KerasModel

        mk_model = lambda seed=None: KerasModel(KerasMLP())  # noqa: E731
    elif a.framework == "flax":
        import jax
        import numpy as np

        from myfyp_b200.learning.frameworks.flax.flax_model import MLP as FlaxMLP
        from myfyp_b200.learning.frameworks.flax.flax_model import FlaxModel

        def mk_model(seed=0):
            module = FlaxMLP()
            return FlaxModel(module, init_params=module.init(jax.random.PRNGKey(seed or 0), np.zeros((1, 28, 28), np.float32))["params"])
    mk_agg = {
        "fedavg": FedAvg, "fedmedian": FedMedian, "trimmedmean": lambda: TrimmedMean(trim=max(1, a.byzantine)), "krum": lambda: Krum(f=a.byzantine),
        "multikrum": lambda: Krum(f=a.byzantine, m=max(1, a.train_set_size - a.byzantine)), "scaffold": lambda: Scaffold(global_lr=1.0), "fedprox": lambda: FedProx(mu=0.01),
    }[a.aggregator]
    proto = InMemoryCommunicationProtocol if a.protocol == "memory" else GrpcCommunicationProtocol

    nodes = []
    for i in range(a.nodes):
        addr = f"node-{i}" if a.protocol == "memory" else (f"unix:///tmp/myfyp-b200-{os.getpid()}-{i}.sock" if a.protocol == "unix" else "127.0.0.1")
        nodes.append(Node(mk_model(seed=a.seed + i), parts[i], address=addr, aggregator=mk_agg(), protocol=proto, batch_size=a.batch_size))
    for n in nodes:
        n.start()
    for n in nodes[a.nodes - a.byzantine :]:
        make_byzantine(n, a.attack, scale=5.0, std=1.0)
    TopologyFactory.connect_nodes(TopologyFactory.generate_matrix(a.topology, a.nodes, seed=a.seed), nodes)
    wait_convergence(nodes, a.nodes - 1, only_direct=False, wait=60)
    fi = FaultInjector()
    if a.drop_node_at_round >= 0:
        fi.drop_at(nodes[-1], a.drop_node_at_round)

    prof = None
    if a.profiling:
        import cProfile

        prof = cProfile.Profile()
        prof.enable()
    t0 = time.time()
    nodes[0].set_start_learning(rounds=a.rounds, epochs=a.epochs)
    alive = nodes[:-1] if a.drop_node_at_round >= 0 else nodes
    wait_to_finish(alive, timeout=3600)
    dt = time.time() - t0
    if prof is not None:
        import pstats

        prof.disable()
        pstats.Stats(prof).sort_stats("cumulative").print_stats(25)
    if a.plot:
        _plot(logger.get_global_logs(), a.plot)
    if a.measure_time:
        print(f"--- {dt:.2f} seconds --- ({a.rounds / dt:.2f} rounds/s)")
    if a.show_metrics:
        glob, local = logger.get_global_logs(), logger.get_local_logs()
        for exp, per_node in glob.items():
            for node, metrics in per_node.items():
                print(exp, node, {k: [round(v, 4) for _, v in series] for k, series in metrics.items()})
        with open(a.out, "w") as f:
            json.dump({"global": glob, "local": {e: {str(r): v for r, v in rounds.items()} for e, rounds in local.items()}, "seconds": dt}, f)
        print(f"metrics written to {a.out}")
    for n in nodes:
        try:
            n.stop()
        except Exception:  # noqa: BLE001
            pass


if __name__ == "__main__":
    main()
