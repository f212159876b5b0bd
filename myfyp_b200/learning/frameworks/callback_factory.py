"""Reference path ``learning/frameworks/callback_factory.py``."""
from myfyp_b200.learning.callbacks.callback import CallbackFactory  # noqa: F401
