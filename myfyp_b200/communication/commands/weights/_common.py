"""Shared error policy of the weights commands."""

import contextlib

from myfyp_b200.learning.exceptions import DecodingParamsError, ModelNotMatchingError, StaleDescriptorError

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


@contextlib.contextmanager
def bad_model_guard(state, stop_fn, source: str, protocol=None):
    """``with bad_model_guard(...)``: anything that goes wrong while a received model is decoded / adopted / filed is the
    SENDER's fault as far as this node can tell - apply :func:`handle_bad_model` instead of letting the handler die."""
    try:
        yield
    except StaleDescriptorError as e:
        # nobody's fault: the owner has reused the slot since it sent the descriptor (we were slow, or the message was kept for a
        # later stage).  The model is refused; the sender still sees us lacking it and offers its current snapshot again.
        logger.warning(state.addr, f"Refused a model from {source}: {e}")
    except DecodingParamsError:
        handle_bad_model(state, stop_fn, source, "Error decoding parameters", protocol)
    except ModelNotMatchingError:
        handle_bad_model(state, stop_fn, source, "Models not matching", protocol)
    except Exception as e:  # noqa: BLE001
        handle_bad_model(state, stop_fn, source, f"Unknown error adding model: {e}", protocol)
