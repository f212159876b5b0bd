"""Shared error handling for weights commands."""

from myfyp_b200.management.logger import logger
from myfyp_b200.settings import Settings


def handle_bad_model(state, stop_fn, source: str, what: str, protocol=None) -> None:
    """A peer sent something we cannot use.

    Reference behaviour is to stop *ourselves* (a DoS vector its own comments admit,
    partial_model_command.py:98-109).  Default here: drop the message and evict the
    sender; ``Settings.ON_BAD_MODEL = "stop"`` restores the reference behaviour.
    """
    logger.error(state.addr, f"{what} (from {source}).")
    if Settings.ON_BAD_MODEL == "stop":
        stop_fn()
    elif protocol is not None:
        try:
            protocol.disconnect(source, disconnect_msg=False)
        except Exception:
            pass
