"""myfyp_b200 - a Blackwell-native decentralized federated-learning framework.

Same capabilities and public API as P2PFL 0.4.0 (the library vendored in PrivEimantas/myFYP): ``Node(model,
data, protocol)``, ``node.connect``, ``set_start_learning(rounds, epochs)``, ``P2PFLModel`` / ``P2PFLDataset``,
the vote -> train -> gossip-aggregate stage machine, FedAvg / FedMedian / Krum / Scaffold (+ FedProx), the
in-memory and gRPC protocols, the encoded-parameter wire/checkpoint format.  The data plane is new: models live
in flat HBM slabs, weights messages carry slab descriptors, and aggregation / diffusion / local training run
as hand-written sm_100a kernels (see ``csrc/`` and ``include/myfyp_b200.h``).
"""

__version__ = "0.1.0"

from myfyp_b200.settings import Settings  # noqa: E402,F401
