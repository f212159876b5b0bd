"""name -> stage class, imported lazily to avoid cycles (reference: p2pfl/stages/stage_factory.py:26-59)."""

from __future__ import annotations

from typing import Type

from myfyp_b200.stages.stage import Stage


class StageFactory:
    @staticmethod
    def get_stage(stage_name: str) -> Type[Stage]:
        if stage_name == "StartLearningStage":
            from myfyp_b200.stages.base_node.start_learning_stage import StartLearningStage as S
        elif stage_name == "VoteTrainSetStage":
            from myfyp_b200.stages.base_node.vote_train_set_stage import VoteTrainSetStage as S
        elif stage_name == "TrainStage":
            from myfyp_b200.stages.base_node.train_stage import TrainStage as S
        elif stage_name == "WaitAggregatedModelsStage":
            from myfyp_b200.stages.base_node.wait_agg_models_stage import WaitAggregatedModelsStage as S
        elif stage_name == "GossipModelStage":
            from myfyp_b200.stages.base_node.gossip_model_stage import GossipModelStage as S
        elif stage_name == "RoundFinishedStage":
            from myfyp_b200.stages.base_node.round_finished_stage import RoundFinishedStage as S
        else:
            raise Exception(f"Invalid stage name: {stage_name}")
        return S
