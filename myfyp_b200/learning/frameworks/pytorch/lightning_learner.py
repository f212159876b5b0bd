"""Reference path ``pytorch/lightning_learner.py``."""
from myfyp_b200.learning.torch_learner import TorchLearner  # noqa: F401

LightningLearner = TorchLearner
