"""Reference path ``learning/frameworks/p2pfl_model.py``."""
from myfyp_b200.learning.p2pfl_model import P2PFLModel  # noqa: F401
