"""Two-process quick start, first half: start a node and wait (reference: p2pfl/examples/node1.py).

    python -m myfyp_b200.examples.node1 --port 6666        then        python -m myfyp_b200.examples.node2 --port 6667 --connect 127.0.0.1:6666
"""

from __future__ import annotations

import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))


def main() -> None:
    ap = argparse.ArgumentParser(description=__doc__.split("\n\n")[0])
    ap.add_argument("--port", type=int, default=6666)
    a = ap.parse_args()
    from myfyp_b200.learning.dataset import P2PFLDataset
    from myfyp_b200.models import mlp_model
    from myfyp_b200.node import Node
    from myfyp_b200.settings import Settings

    Settings.USE_SSL = False
    node = Node(mlp_model(seed=1), P2PFLDataset.synthetic_mnist(600, 100, seed=1), address=f"127.0.0.1:{a.port}")
    print(f"node listening on {node.addr}; Ctrl-C to stop")
    try:
        node.start(wait=True)
    except KeyboardInterrupt:
        node.stop()


if __name__ == "__main__":
    main()
