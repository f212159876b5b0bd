"""Protocol exceptions (reference: communication/protocols/exceptions.py:21-36)."""


class ProtocolNotStartedError(Exception):
    """The protocol is used before ``start()`` / after ``stop()``."""


class NeighborNotConnectedError(Exception):
    """Send to an address that is not a (direct) neighbour."""


def peer_looks_dead(e: BaseException) -> bool:
    """Did this send fail because the PEER is gone (its server does not run, the connection is refused / broken / silent) - as
    opposed to the peer answering with an error of its own (a handler that raised, an unknown command)?  Only the first kind is
    worth telling the rest of the network (``node_down``); the second says nothing about the peer's health."""
    if isinstance(e, (ConnectionError, FileNotFoundError, TimeoutError, EOFError, BrokenPipeError, OSError)):
        return True
    text = str(e)
    if isinstance(e, CommunicationError):
        return "server not running" in text
    if isinstance(e, NeighborNotConnectedError):
        return False
    return type(e).__name__ in ("RpcError", "_InactiveRpcError", "_MultiThreadedRendezvous") or "is gone" in text


class CommunicationError(Exception):
    """The peer answered with an error (or could not be reached)."""
