"""Reference path ``learning/frameworks/learner_factory.py``."""
from myfyp_b200.learning.learner import LearnerFactory  # noqa: F401
