"""VoteTrainSetStage (reference: stages/base_node/vote_train_set_stage.py:43-184).

Vote: sample ``min(TRAIN_SET_SIZE, n)`` candidates among all known nodes + self with weights
``floor(randint(0,1000)/(i+1))``, broadcast.  Tally when every live neighbour has voted or at
``VOTE_TIMEOUT``: sum weights, order by address descending then (stable) by votes descending, keep the top
``TRAIN_SET_SIZE`` (SURVEY Q18 - every node must reach the same set), drop nodes that vanished.

Event-driven: instead of re-polling every 2 s the stage sleeps on the node state's condition variable, which
vote arrivals and neighbour-table changes notify.
"""

from __future__ import annotations

import math
import random
import time
from typing import Dict, List, Type, Union

from myfyp_b200.communication.commands.message.vote_train_set_command import VoteTrainSetCommand
from myfyp_b200.management.logger import logger
from myfyp_b200.settings import Settings
from myfyp_b200.stages.stage import EarlyStopException, Stage, check_early_stop
from myfyp_b200.stages.stage_factory import StageFactory


class VoteTrainSetStage(Stage):
    @staticmethod
    def name() -> str:
        return "VoteTrainSetStage"

    @staticmethod
    def execute(state=None, communication_protocol=None, **kwargs) -> Union[Type[Stage], None]:
        if state is None or communication_protocol is None:
            raise Exception("Invalid parameters on VoteTrainSetStage.")
        logger.round_started(state.addr, state.experiment)
        try:
            VoteTrainSetStage.__vote(state, communication_protocol)
            state.train_set = VoteTrainSetStage.__validate_train_set(VoteTrainSetStage.__aggregate_votes(state, communication_protocol), state, communication_protocol)
            logger.info(state.addr, f"🚂 Train set of {len(state.train_set)} nodes: {state.train_set}")
            if state.addr in state.train_set:
                return StageFactory.get_stage("TrainStage")
            return StageFactory.get_stage("WaitAggregatedModelsStage")
        except EarlyStopException:
            return None

    @staticmethod
    def __vote(state, communication_protocol) -> None:
        candidates = list(communication_protocol.get_neighbors(only_direct=False))
        if state.addr not in candidates:
            candidates.append(state.addr)
        samples = min(Settings.TRAIN_SET_SIZE, len(candidates))
        voted = random.sample(candidates, samples)
        weights = [math.floor(random.randint(0, 1000) / (i + 1)) for i in range(samples)]
        votes = list(zip(voted, weights))
        with state.train_set_votes_lock:
            state.train_set_votes[state.addr] = dict(votes)
        logger.info(state.addr, "🗳️ Sending train set vote.")
        flat = [str(x) for pair in votes for x in pair]
        communication_protocol.broadcast(communication_protocol.build_msg(VoteTrainSetCommand.get_name(), flat, round=state.round))

    @staticmethod
    def __aggregate_votes(state, communication_protocol) -> List[str]:
        deadline = time.monotonic() + Settings.VOTE_TIMEOUT
        while True:
            check_early_stop(state)
            neis = set(communication_protocol.get_neighbors(only_direct=False))
            needed = neis | {state.addr}
            # Every ballot received for this round counts, also the ballot of a node that has left the neighbour table since it
            # voted: whether a dead node's ballot is in the tally must not depend on WHEN a node noticed the death, or the
            # survivors elect different train sets.  (The reference filters by the current neighbours, :137-141.)
            with state.train_set_votes_lock:
                votes = dict(state.train_set_votes)
            ready = needed <= set(votes.keys())
            timeout = time.monotonic() > deadline
            if ready or timeout:
                if timeout and not ready:
                    logger.info(state.addr, f"Timeout for vote aggregation. Missing votes from {needed - set(votes.keys())}")
                tally: Dict[str, int] = {}
                for ballot in votes.values():
                    for k, v in ballot.items():
                        tally[k] = tally.get(k, 0) + v
                ordered = sorted(tally.items(), key=lambda kv: kv[0], reverse=True)  # tie-break: address, descending
                ordered = sorted(ordered, key=lambda kv: kv[1], reverse=True)  # stable
                top = ordered[: min(len(ordered), Settings.TRAIN_SET_SIZE)]
                with state.train_set_votes_lock:
                    state.train_set_votes = {}
                logger.info(state.addr, f"🔢 Computed {len(votes)} votes.")
                return [k for k, _ in top]
            # woken by a vote, a neighbour change or stop; the timeout only bounds a lost wake-up
            n_votes = len(votes)
            state.wait_for(lambda: state.round is None or len(state.train_set_votes) != n_votes, timeout=min(0.5, max(deadline - time.monotonic(), 0.01)))

    @staticmethod
    def __validate_train_set(train_set: List[str], state, communication_protocol) -> List[str]:
        alive = set(communication_protocol.get_neighbors(only_direct=False))
        return [n for n in train_set if n in alive or n == state.addr]
