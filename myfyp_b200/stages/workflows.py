"""Stage workflows (reference: p2pfl/stages/workflows.py:28-60).

``history`` and ``finished`` are part of the de-facto public API (tests assert on both).  Per-stage wall
times are kept in ``timings`` for the rounds/s breakdown (B200 addition).
"""

from __future__ import annotations

import time
from typing import List, Tuple, Type

from myfyp_b200.stages.stage import Stage
from myfyp_b200.stages.stage_factory import StageFactory


from myfyp_b200.management.nvtx import nvtx_range


class StageWorkflow:
    def __init__(self, first_stage: Type[Stage]) -> None:
        self.current_stage = first_stage
        self.history: List[str] = []
        self.timings: List[Tuple[str, float]] = []
        self.finished = False

    def run(self, **kwargs) -> None:
        self.finished = False
        while True:
            name = self.current_stage.name()
            self.history.append(name)
            t0 = time.perf_counter()
            st = kwargs.get("state")
            with nvtx_range(f"{getattr(st, 'addr', '?')}/{name}"):
                next_stage = self.current_stage.execute(**kwargs)
            self.timings.append((name, time.perf_counter() - t0))
            if next_stage is None:  # every stage turns an early stop into None itself
                self.finished = True
                break
            self.current_stage = next_stage


# the reference spells it "StageWokflow"; keep the alias for scripts that import it
StageWokflow = StageWorkflow


class LearningWorkflow(StageWorkflow):
    def __init__(self) -> None:
        super().__init__(StageFactory.get_stage("StartLearningStage"))

    def run(self, **kwargs) -> None:
        self.current_stage = StageFactory.get_stage(kwargs.pop("first_stage", "StartLearningStage"))
        super().run(**kwargs)
