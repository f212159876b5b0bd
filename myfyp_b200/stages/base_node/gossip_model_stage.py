"""GossipModelStage (reference: stages/base_node/gossip_model_stage.py:41-87): push the full aggregated model
(``add_model``) to direct neighbours whose announced round is behind ours, until none is left."""

from __future__ import annotations

from typing import List, Type, Union

from myfyp_b200.communication.commands.weights.full_model_command import FullModelCommand
from myfyp_b200.management.logger import logger
from myfyp_b200.stages.diffusion import Diffusion
from myfyp_b200.stages.stage import Stage
from myfyp_b200.stages.stage_factory import StageFactory


class GossipModelStage(Stage):
    @staticmethod
    def name() -> str:
        return "GossipModelStage"

    @staticmethod
    def execute(state=None, communication_protocol=None, aggregator=None, learner=None, **kwargs) -> Union[Type[Stage], None]:
        if None in (state, aggregator, communication_protocol, learner):
            raise Exception("Invalid parameters on GossipModelStage.")
        this_round = state.round
        if this_round is None:
            return None
        logger.info(state.addr, "🗣️ Diffusing the aggregated model.")

        def behind() -> List[str]:
            # a neighbour we never heard a status from is behind by definition (the reference raises KeyError there)
            return [n for n in communication_protocol.get_neighbors(only_direct=True) if state.nei_status.get(n, -1) < this_round]

        # one snapshot / one pickle per stage (cached by the diffusion), not one per neighbour
        Diffusion(state, communication_protocol, FullModelCommand.get_name(), behind, learner.get_model).run()
        return StageFactory.get_stage("RoundFinishedStage")
