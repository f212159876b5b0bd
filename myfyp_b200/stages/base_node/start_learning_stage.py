"""StartLearningStage (reference: stages/base_node/start_learning_stage.py:44-112).

Install the experiment, wait until this node holds the initial model (the initiator owns it, everybody else blocks until an
``init_model`` arrives - interruptibly, unlike the reference's bare ``lock.acquire()``), announce that, hand the model on to
direct neighbours that have not announced theirs, give fresh heartbeats time to flood, then vote.
"""

from __future__ import annotations

import threading
import time
from typing import Optional, Type, Union

from myfyp_b200.communication.commands.message.model_initialized_command import ModelInitializedCommand
from myfyp_b200.communication.commands.weights.init_model_command import InitModelCommand
from myfyp_b200.management.logger import logger
from myfyp_b200.settings import Settings
from myfyp_b200.stages.diffusion import Diffusion
from myfyp_b200.stages.stage import Stage
from myfyp_b200.stages.stage_factory import StageFactory

def _holds_initial_model(state) -> bool:
    """Block (in slices, so a stop is noticed) until the model-initialised lock can be taken; False when the experiment stopped."""
    while not state.model_initialized_lock.acquire(timeout=0.25):
        if state.round is None:
            return False
    return True


class StartLearningStage(Stage):
    @staticmethod
    def name() -> str:
        return "StartLearningStage"

    @staticmethod
    def execute(rounds: Optional[int] = None, epochs: Optional[int] = None, state=None, learner=None, communication_protocol=None, aggregator=None, start_round: int = 0, **kwargs) -> Union[Type[Stage], None]:
        if None in (rounds, epochs, state, learner, communication_protocol, aggregator):
            raise Exception("Invalid parameters on StartLearningStage.")
        t_start = time.monotonic()
        with state.start_thread_lock:  # one learning thread per node
            state.set_experiment(Settings.EXPERIMENT_NAME, rounds, round=start_round)
            learner.set_epochs(epochs)
            logger.experiment_started(state.addr, state.experiment)
            threading.current_thread().b2_experiment_set = True  # type: ignore[attr-defined]  (Node tells a stopped thread from a duplicate start by it)

        logger.info(state.addr, "⏳ Waiting for the initial model.")
        if not _holds_initial_model(state):
            return None
        communication_protocol.broadcast(communication_protocol.build_msg(ModelInitializedCommand.get_name()))

        logger.info(state.addr, "🗣️ Passing the initial model on.")
        uninitialised = lambda: [n for n in communication_protocol.get_neighbors(only_direct=True) if n not in state.nei_status]  # noqa: E731
        Diffusion(state, communication_protocol, InitModelCommand.get_name(), uninitialised, learner.get_model).run()

        # heartbeats of freshly connected nodes must have flooded before the vote counts candidates
        settle = Settings.WAIT_HEARTBEATS_CONVERGENCE - (time.monotonic() - t_start)
        if settle > 0:
            time.sleep(settle)
        return StageFactory.get_stage("VoteTrainSetStage")
