"""Learning-layer exceptions (reference: learning/frameworks/exceptions.py:22-31)."""


class DecodingParamsError(Exception):
    """A serialized parameter blob could not be decoded."""


class ModelNotMatchingError(Exception):
    """Received parameters do not fit this model's architecture."""


class StaleDescriptorError(Exception):
    """A slab descriptor names an arena slot that its owner has published something newer into since."""
