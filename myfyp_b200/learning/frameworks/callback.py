"""Reference path ``learning/frameworks/callback.py``."""
from myfyp_b200.learning.callbacks.callback import P2PFLCallback  # noqa: F401
