"""Reference path ``pytorch/lightning_model.py``: ``LightningModel`` is the slab-backed ``TorchP2PFLModel``; ``MLP`` the example model."""
from myfyp_b200.learning.p2pfl_model import LightningModel, TorchP2PFLModel  # noqa: F401
from myfyp_b200.models.mlp import MLP  # noqa: F401
