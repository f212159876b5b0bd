"""StartLearningStage (reference: stages/base_node/start_learning_stage.py:44-112).

Set the experiment, wait until this node's model is initialised (the initiator owns the weights, everyone
else blocks until an ``init_model`` arrives), announce it, gossip the initial weights to direct neighbours
that have not announced theirs, let heartbeats converge, then vote.
"""

from __future__ import annotations

import time
from typing import Any, List, Optional, Type, Union

from myfyp_b200.communication.commands.message.model_initialized_command import ModelInitializedCommand
from myfyp_b200.communication.commands.weights.init_model_command import InitModelCommand
from myfyp_b200.management.logger import logger
from myfyp_b200.settings import Settings
from myfyp_b200.stages.stage import Stage
from myfyp_b200.stages.stage_factory import StageFactory


class StartLearningStage(Stage):
    @staticmethod
    def name() -> str:
        return "StartLearningStage"

    @staticmethod
    def execute(rounds: Optional[int] = None, epochs: Optional[int] = None, state=None, learner=None, communication_protocol=None, aggregator=None, start_round: int = 0, **kwargs) -> Union[Type[Stage], None]:
        if rounds is None or epochs is None or state is None or learner is None or communication_protocol is None or aggregator is None:
            raise Exception("Invalid parameters on StartLearningStage.")
        with state.start_thread_lock:
            state.set_experiment(Settings.EXPERIMENT_NAME, rounds, round=start_round)
            learner.set_epochs(epochs)
            logger.experiment_started(state.addr, state.experiment)
        begin = time.monotonic()

        logger.info(state.addr, "⏳ Waiting initialization.")
        # interruptible version of the reference's blocking lock.acquire()
        while not state.model_initialized_lock.acquire(timeout=0.25):
            if state.round is None:
                return None
        communication_protocol.broadcast(communication_protocol.build_msg(ModelInitializedCommand.get_name()))

        logger.info(state.addr, "🗣️ Gossiping model initialization.")
        StartLearningStage.__gossip_model(state, communication_protocol, learner)

        # heartbeats of freshly connected nodes must have flooded before the vote counts candidates
        left = Settings.WAIT_HEARTBEATS_CONVERGENCE - (time.monotonic() - begin)
        if left > 0:
            time.sleep(left)
        return StageFactory.get_stage("VoteTrainSetStage")

    @staticmethod
    def __gossip_model(state, communication_protocol, learner) -> None:
        def early_stopping_fn() -> bool:
            return state.round is None

        def get_candidates_fn() -> List[str]:
            return [n for n in communication_protocol.get_neighbors(only_direct=True) if n not in state.nei_status]

        def model_fn(nei: str) -> Any:
            if state.round is None:
                raise Exception("Round not initialized.")
            model = learner.get_model()
            payload = _transport(model, communication_protocol, nei)
            return communication_protocol.build_weights(InitModelCommand.get_name(), state.round, payload)

        communication_protocol.gossip_weights(
            early_stopping_fn, get_candidates_fn, get_candidates_fn, model_fn, wait_fn=state.wait_change
        )


def _transport(model, communication_protocol, nei: str):
    """Device descriptor when ``nei`` can dereference it, the reference pickle otherwise."""
    enc = getattr(model, "encode_for_transport", None)
    if enc is None:
        return model.encode_parameters()
    return enc(communication_protocol.fabric_reachable(nei))
