"""Learning-layer exceptions (reference: learning/frameworks/exceptions.py:22-31)."""


class DecodingParamsError(Exception):
    """A serialized parameter blob could not be decoded."""


class ModelNotMatchingError(Exception):
    """Received parameters do not fit this model's architecture."""
