"""Reference path ``learning/frameworks/learner.py``."""
from myfyp_b200.learning.learner import Learner  # noqa: F401
