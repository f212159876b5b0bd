"""Reference path ``learning/frameworks/exceptions.py``."""
from myfyp_b200.learning.exceptions import DecodingParamsError, ModelNotMatchingError  # noqa: F401
