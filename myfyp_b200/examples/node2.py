"""Two-process quick start, second half: connect to node1 and start learning (reference: p2pfl/examples/node2.py)."""

from __future__ import annotations

import argparse
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))


def main() -> None:
    ap = argparse.ArgumentParser(description=__doc__.split("\n\n")[0])
    ap.add_argument("--port", type=int, default=6667)
    ap.add_argument("--connect", default="127.0.0.1:6666")
    ap.add_argument("--rounds", type=int, default=2)
    ap.add_argument("--epochs", type=int, default=1)
    a = ap.parse_args()
    from myfyp_b200.learning.dataset import P2PFLDataset
    from myfyp_b200.management.logger import logger
    from myfyp_b200.models import mlp_model
    from myfyp_b200.node import Node
    from myfyp_b200.settings import Settings
    from myfyp_b200.utils.utils import wait_to_finish

    Settings.USE_SSL = False
    node = Node(mlp_model(seed=2), P2PFLDataset.synthetic_mnist(600, 100, seed=2), address=f"127.0.0.1:{a.port}")
    node.start()
    if not node.connect(a.connect):
        raise SystemExit(f"cannot reach {a.connect}")
    time.sleep(2 * Settings.HEARTBEAT_PERIOD)
    node.set_start_learning(rounds=a.rounds, epochs=a.epochs)
    wait_to_finish([node], timeout=600)
    print(logger.get_global_logs())
    node.stop()


if __name__ == "__main__":
    main()
