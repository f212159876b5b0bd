"""Protocol exceptions (reference: communication/protocols/exceptions.py:21-36)."""


class ProtocolNotStartedError(Exception):
    """The protocol is used before ``start()`` / after ``stop()``."""


class NeighborNotConnectedError(Exception):
    """Send to an address that is not a (direct) neighbour."""


class CommunicationError(Exception):
    """The peer answered with an error (or could not be reached)."""
