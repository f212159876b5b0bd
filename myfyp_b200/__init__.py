"""myfyp_b200 - a Blackwell-native decentralized federated-learning framework.

Same capabilities and public API as P2PFL 0.4.0 (the library vendored in PrivEimantas/myFYP): ``Node(model,
data, protocol)``, ``node.connect``, ``set_start_learning(rounds, epochs)``, ``P2PFLModel`` / ``P2PFLDataset``,
the vote -> train -> gossip-aggregate stage machine, FedAvg / FedMedian / Krum / Scaffold (+ FedProx), the
in-memory and gRPC protocols, the encoded-parameter wire/checkpoint format.  The data plane is new: models live
in flat HBM slabs, weights messages carry slab descriptors, and aggregation / diffusion / local training run
as hand-written sm_100a kernels (see ``csrc/`` and ``include/myfyp_b200.h``).
"""

__version__ = "0.1.0"

import os as _os

# Every node trains on its own CUDA stream (plus a side stream for the weight-gradient branch) and several nodes share a GPU.
# CUDA maps streams onto 8 hardware work queues by default; with more active streams than queues, unrelated streams share a
# queue and one node's whole epoch graph falsely serialises behind another's (measured: 4 nodes on one B200 finished their
# fits in two waves, 49 ms instead of 35 ms).  32 is the maximum.  Must be set before the CUDA context exists, hence here.
_os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")

from myfyp_b200.settings import Settings  # noqa: E402,F401
