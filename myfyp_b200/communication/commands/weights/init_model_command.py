"""``init_model``: the initiator's starting weights, handed from initialised nodes to their uninitialised neighbours
(reference: commands/weights/init_model_command.py:31-97).  Adopted at most once: the node's ``model_initialized_lock`` is
created held and released by the first delivery; concurrent deliveries are serialised by a per-command mutex (in the reference
two of them both release the lock and the second one stops the node)."""

import threading
from typing import Callable

from myfyp_b200.communication.commands.command import Command
from myfyp_b200.communication.commands.weights._common import bad_model_guard
from myfyp_b200.management.logger import logger


class InitModelCommand(Command):
    def __init__(self, state, stop: Callable[[], None], aggregator, learner, protocol=None) -> None:
        self.state, self.stop, self.aggregator, self.learner, self.protocol = state, stop, aggregator, learner, protocol
        self._mutex = threading.Lock()

    @staticmethod
    def get_name() -> str:
        return "init_model"

    def _why_not(self, round: int, weights) -> str:
        st = self.state
        if weights is None:
            return "no weights in the message"
        if st.round is None:
            return "learning is not running"
        if round != st.round:
            return f"it belongs to round {round}, this node is in round {st.round}"
        return ""

    def execute(self, source: str, round: int, weights=None, **kwargs) -> None:
        st = self.state
        reason = self._why_not(round, weights)
        if reason:
            (logger.error if weights is None else logger.debug)(st.addr, f"init_model from {source} ignored: {reason}.")
            return
        with bad_model_guard(st, self.stop, source, self.protocol):
            with self._mutex:
                if not st.model_initialized_lock.locked():
                    logger.debug(st.addr, f"init_model from {source} ignored: the model is initialised already.")
                    return
                self.learner.set_model(weights)
                st.model_initialized_lock.release()
            st.notify()
            logger.info(st.addr, "🤖 Initial model adopted.")
