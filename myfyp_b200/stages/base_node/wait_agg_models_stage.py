"""WaitAggregatedModelsStage (reference: stages/base_node/wait_agg_models_stage.py:40-67): nodes outside
the train set wait (<= AGGREGATION_TIMEOUT) for the full aggregated model, then announce ``models_ready``."""

from __future__ import annotations

from typing import Type, Union

from myfyp_b200.communication.commands.message.models_ready_command import ModelsReadyCommand
from myfyp_b200.management.logger import logger
from myfyp_b200.settings import Settings
from myfyp_b200.stages.stage import Stage
from myfyp_b200.stages.stage_factory import StageFactory


class WaitAggregatedModelsStage(Stage):
    @staticmethod
    def name() -> str:
        return "WaitAggregatedModelsStage"

    @staticmethod
    def execute(state=None, communication_protocol=None, **kwargs) -> Union[Type[Stage], None]:
        if state is None or communication_protocol is None:
            raise Exception("Invalid parameters on WaitAggregatedModelsStage.")
        state.aggregated_model_event.clear()
        logger.info(state.addr, "⏳ Waiting aggregation.")
        ok = state.wait_for(lambda: state.aggregated_model_event.is_set() or state.round is None, timeout=Settings.AGGREGATION_TIMEOUT)
        if state.round is None:
            return None
        if ok:
            logger.info(state.addr, "✅ Aggregation event received.")
        else:
            logger.warning(state.addr, "⏰ Aggregation timeout occurred.")
            state.aggregated_model_event.set()
        communication_protocol.broadcast(communication_protocol.build_msg(ModelsReadyCommand.get_name(), [], round=state.round))
        return StageFactory.get_stage("GossipModelStage")
