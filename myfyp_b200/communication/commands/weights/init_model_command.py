"""``init_model`` (reference: commands/weights/init_model_command.py:31-97)."""

import threading
from typing import Callable

from myfyp_b200.communication.commands.command import Command
from myfyp_b200.communication.commands.weights._common import handle_bad_model
from myfyp_b200.learning.exceptions import DecodingParamsError, ModelNotMatchingError
from myfyp_b200.management.logger import logger


class InitModelCommand(Command):
    def __init__(self, state, stop: Callable[[], None], aggregator, learner, protocol=None) -> None:
        self.state, self.stop, self.aggregator, self.learner, self.protocol = state, stop, aggregator, learner, protocol
        self._mutex = threading.Lock()  # several initialised neighbours may push init_model at the same instant

    @staticmethod
    def get_name() -> str:
        return "init_model"

    def execute(self, source: str, round: int, weights=None, **kwargs) -> None:
        st = self.state
        if weights is None:
            logger.error(st.addr, "Invalid InitModelCommand message")
            return
        if st.round is None:
            logger.debug(st.addr, "Tried to initialise a model while learning is not running")
            return
        if round != st.round:
            logger.debug(st.addr, f"Model initialisation in a late round ({round} != {st.round}).")
            return
        try:
            with self._mutex:
                if not st.model_initialized_lock.locked():
                    logger.debug(st.addr, "Model initialisation message when the model is already initialised. Ignored.")
                    return
                self.learner.set_model(weights)
                st.model_initialized_lock.release()
            st.notify()
            logger.info(st.addr, "🤖 Model Weights Initialized")
        except DecodingParamsError:
            handle_bad_model(st, self.stop, source, "Error decoding parameters", self.protocol)
        except ModelNotMatchingError:
            handle_bad_model(st, self.stop, source, "Models not matching", self.protocol)
        except Exception as e:  # noqa: BLE001
            handle_bad_model(st, self.stop, source, f"Unknown error adding model: {e}", self.protocol)
