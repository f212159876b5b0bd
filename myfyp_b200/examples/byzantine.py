"""Robust aggregation under model poisoning: FedAvg vs FedMedian vs Krum with sign-flip / noise attackers.

The fork author's experiment (exp_SAVE3.txt: one of three nodes in a star flips the sign of, or adds Gaussian noise to,
its weights) as a reproducible script: prints the honest nodes' final test accuracy per aggregator.
"""

from __future__ import annotations

import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))


def main() -> None:
    ap = argparse.ArgumentParser(description=__doc__.split("\n\n")[0])
    ap.add_argument("--nodes", type=int, default=5)
    ap.add_argument("--byzantine", type=int, default=1)
    ap.add_argument("--attack", default="sign_flip", choices=["sign_flip", "gaussian", "scaled_update"])
    ap.add_argument("--rounds", type=int, default=3)
    a = ap.parse_args()
    import numpy as np

    from myfyp_b200.communication.protocols.memory.memory_communication_protocol import InMemoryCommunicationProtocol
    from myfyp_b200.learning.aggregators import FedAvg, FedMedian, Krum
    from myfyp_b200.learning.dataset import P2PFLDataset, RandomIIDPartitionStrategy
    from myfyp_b200.management.logger import logger
    from myfyp_b200.models import mlp_model
    from myfyp_b200.node import Node
    from myfyp_b200.settings import Settings
    from myfyp_b200.utils.fault_injector import make_byzantine
    from myfyp_b200.utils.topologies import TopologyFactory
    from myfyp_b200.utils.utils import set_standalone_settings, wait_convergence, wait_to_finish

    set_standalone_settings()
    logger.set_level("WARNING")
    Settings.TRAIN_SET_SIZE = a.nodes
    parts = P2PFLDataset.synthetic_mnist(300 * a.nodes, 60 * a.nodes, seed=7).generate_partitions(a.nodes, RandomIIDPartitionStrategy)
    for name, mk in (("FedAvg", FedAvg), ("FedMedian", FedMedian), (f"Krum(f={a.byzantine})", lambda: Krum(f=a.byzantine, m=a.nodes - 2 * a.byzantine))):
        logger.reset_metrics()
        nodes = [Node(mlp_model(seed=i), parts[i], address=f"{name}-{i}", aggregator=mk(), protocol=InMemoryCommunicationProtocol) for i in range(a.nodes)]
        for n in nodes:
            n.start()
        for n in nodes[a.nodes - a.byzantine :]:
            make_byzantine(n, a.attack, scale=5.0)
        TopologyFactory.connect_nodes(TopologyFactory.generate_matrix("star", a.nodes), nodes)
        wait_convergence(nodes, a.nodes - 1, only_direct=False, wait=30)
        nodes[0].set_start_learning(rounds=a.rounds, epochs=1)
        wait_to_finish(nodes, timeout=600)
        glob = logger.get_global_logs()[Settings.EXPERIMENT_NAME]
        acc = float(np.mean([glob[n.addr]["test_metric"][-1][1] for n in nodes[: a.nodes - a.byzantine]]))
        print(f"{name:14s} honest-node accuracy after {a.rounds} rounds with {a.byzantine} {a.attack} attacker(s): {acc:.3f}")
        for n in nodes:
            n.stop()


if __name__ == "__main__":
    main()
