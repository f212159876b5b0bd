"""RoundFinishedStage (reference: stages/base_node/round_finished_stage.py:42-91)."""

from __future__ import annotations

from typing import Type, Union

from myfyp_b200.management.logger import logger
from myfyp_b200.settings import Settings
from myfyp_b200.stages.base_node.train_stage import evaluate_and_broadcast
from myfyp_b200.stages.stage import Stage
from myfyp_b200.stages.stage_factory import StageFactory


class RoundFinishedStage(Stage):
    @staticmethod
    def name() -> str:
        return "RoundFinishedStage"

    @staticmethod
    def execute(state=None, learner=None, communication_protocol=None, aggregator=None, **kwargs) -> Union[Type[Stage], None]:
        if state is None or communication_protocol is None or aggregator is None or learner is None:
            raise Exception("Invalid parameters on RoundFinishedStage.")
        if state.round is None:
            return None
        aggregator.clear()
        state.increase_round()
        logger.round_finished(state.addr)
        logger.info(state.addr, f"🎉 Round {state.round} of {state.total_rounds} finished.")
        if Settings.CHECKPOINT_DIR:
            # the model every node holds after the round, in the wire layout, plus the experiment position (resume: utils/checkpoint.py)
            try:
                from myfyp_b200.utils.checkpoint import save_checkpoint

                name = state.addr.replace(":", "_").replace("/", "_")
                save_checkpoint(f"{Settings.CHECKPOINT_DIR}/{name}.ckpt", learner.get_model(), state.round, state.total_rounds, state.exp_name)
            except Exception as e:  # noqa: BLE001 - a full disk must not stop the federation
                logger.error(state.addr, f"checkpoint failed: {e}")
        if state.round is None or state.total_rounds is None:
            raise ValueError("Round or total rounds not set.")
        if state.round < state.total_rounds:
            return StageFactory.get_stage("VoteTrainSetStage")
        evaluate_and_broadcast(state, learner, communication_protocol)
        state.clear()
        logger.info(state.addr, "😋 Training finished!!")
        return None
