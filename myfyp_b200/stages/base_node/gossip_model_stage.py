"""GossipModelStage (reference: stages/base_node/gossip_model_stage.py:41-87): push the full aggregated model
(``add_model``) to direct neighbours whose announced round is behind ours, until none is left."""

from __future__ import annotations

from typing import Any, List, Type, Union

from myfyp_b200.communication.commands.weights.full_model_command import FullModelCommand
from myfyp_b200.management.logger import logger
from myfyp_b200.stages.base_node.start_learning_stage import _transport
from myfyp_b200.stages.stage import Stage, check_early_stop
from myfyp_b200.stages.stage_factory import StageFactory


class GossipModelStage(Stage):
    @staticmethod
    def name() -> str:
        return "GossipModelStage"

    @staticmethod
    def execute(state=None, communication_protocol=None, aggregator=None, learner=None, **kwargs) -> Union[Type[Stage], None]:
        if state is None or aggregator is None or communication_protocol is None or learner is None:
            raise Exception("Invalid parameters on GossipModelStage.")
        if state.round is None:
            return None
        logger.info(state.addr, "🗣️ Gossiping aggregated model.")
        fixed_round = state.round
        published = {}

        def get_candidates_fn() -> List[str]:
            # a neighbour we never heard a status from is behind by definition (reference raises KeyError here)
            return [n for n in communication_protocol.get_neighbors(only_direct=True) if state.nei_status.get(n, -1) < fixed_round]

        def model_fn(node: str) -> Any:
            if state.round is None:
                raise Exception("Round not initialized")
            reach = communication_protocol.fabric_reachable(node)
            if reach not in published:  # one snapshot / one pickle per stage, not one per neighbour
                published[reach] = _transport(learner.get_model(), communication_protocol, node)
            return communication_protocol.build_weights(FullModelCommand.get_name(), state.round, published[reach])

        communication_protocol.gossip_weights(
            lambda: check_early_stop(state, raise_exception=False), get_candidates_fn, get_candidates_fn, model_fn, wait_fn=state.wait_change
        )
        return StageFactory.get_stage("RoundFinishedStage")
