"""Reference path ``pytorch/lightning_dataset.py``."""
from myfyp_b200.learning.dataset.p2pfl_dataset import PyTorchExportStrategy, TensorExportStrategy  # noqa: F401
