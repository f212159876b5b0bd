"""WaitAggregatedModelsStage (reference: stages/base_node/wait_agg_models_stage.py:40-67): a node outside the round's train
set waits - on the node's condition variable, at most AGGREGATION_TIMEOUT - until a ``FullModelCommand`` has delivered the
aggregated model, then tells everybody that it is ready for the next round."""

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
        arrived = state.aggregated_model_event
        arrived.clear()
        logger.info(state.addr, "⏳ Waiting for the aggregated model.")
        if getattr(state, "adopt_early_full_model", None) is not None:
            state.adopt_early_full_model()  # a neighbour may have pushed it while this node was still tallying votes
        in_time = state.wait_for(lambda: arrived.is_set() or state.round is None, timeout=Settings.AGGREGATION_TIMEOUT)
        if state.round is None:  # stopped while waiting
            return None
        if not in_time:
            logger.warning(state.addr, f"⏰ No aggregated model within {Settings.AGGREGATION_TIMEOUT} s: going on with the local one.")
            arrived.set()
        communication_protocol.broadcast(communication_protocol.build_msg(ModelsReadyCommand.get_name(), [], round=state.round))
        return StageFactory.get_stage("GossipModelStage")
