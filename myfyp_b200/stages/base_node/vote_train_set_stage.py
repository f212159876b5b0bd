"""VoteTrainSetStage (reference: stages/base_node/vote_train_set_stage.py:43-184).

Election of the round's train set, in three steps every node performs on its own:

1. ``cast_ballot``   pick ``min(TRAIN_SET_SIZE, n)`` candidates among all known nodes + self, weight the i-th pick
                     ``floor(randint(0, 1000) / (i + 1))``, flood the ballot;
2. ``collect``       sleep on the node's condition variable until a ballot of every live neighbour is in (ballot arrivals and
                     neighbour-table changes wake it) or ``VOTE_TIMEOUT`` passes;
3. ``elect``         add up ALL ballots received for this round - also those of nodes that have left the neighbour table since
                     (otherwise survivors of a node death elect different sets) - rank by votes, ties by address (descending,
                     SURVEY Q18), keep the first ``TRAIN_SET_SIZE`` that are still alive.
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

Ballot = Dict[str, int]


def cast_ballot(state, communication_protocol) -> Ballot:
    known = set(communication_protocol.get_neighbors(only_direct=False)) | {state.addr}
    picks = random.sample(sorted(known), min(Settings.TRAIN_SET_SIZE, len(known)))
    ballot: Ballot = {node: math.floor(random.randint(0, 1000) / (rank + 1)) for rank, node in enumerate(picks)}
    with state.train_set_votes_lock:
        state.train_set_votes[state.addr] = dict(ballot)
    logger.info(state.addr, "🗳️ Ballot cast.")
    wire = [str(x) for pair in ballot.items() for x in pair]
    communication_protocol.broadcast(communication_protocol.build_msg(VoteTrainSetCommand.get_name(), wire, round=state.round))
    return ballot


def collect(state, communication_protocol) -> Dict[str, Ballot]:
    """Every ballot of the round once all live neighbours have voted (or at VOTE_TIMEOUT); clears the ballot box."""
    deadline = time.monotonic() + Settings.VOTE_TIMEOUT
    while True:
        check_early_stop(state)
        v0 = state.version()
        electorate = set(communication_protocol.get_neighbors(only_direct=False)) | {state.addr}
        with state.train_set_votes_lock:
            box = dict(state.train_set_votes)
        missing = electorate - set(box)
        timed_out = time.monotonic() > deadline
        if not missing or timed_out:
            if missing:
                logger.info(state.addr, f"Vote time-out: no ballot from {sorted(missing)}")
            with state.train_set_votes_lock:
                state.train_set_votes = {}
            return box
        # woken by the ballot that completes the box (VoteTrainSetCommand reads votes_needed), a neighbour change or a stop; the
        # short time-out only bounds a lost wake-up
        # (a ballot that arrives before votes_needed is set notifies unconditionally and bumps the version read above, so it is not
        # missed; afterwards only ballots that can complete the box wake this thread)
        state.votes_needed = len(electorate)
        try:
            state.wait_since(v0, timeout=min(0.5, max(deadline - time.monotonic(), 0.01)))
        finally:
            state.votes_needed = None


def elect(box: Dict[str, Ballot], alive: set, size: int) -> List[str]:
    totals: Dict[str, int] = {}
    for ballot in box.values():
        for node, weight in ballot.items():
            totals[node] = totals.get(node, 0) + weight
    # most votes first; equal votes: larger address first - every node must order ties the same way
    ranking = sorted(sorted(totals, reverse=True), key=lambda n: totals[n], reverse=True)
    return [n for n in ranking[:size] if n in alive]


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
            cast_ballot(state, communication_protocol)
            box = collect(state, communication_protocol)
            alive = set(communication_protocol.get_neighbors(only_direct=False)) | {state.addr}
            state.train_set = elect(box, alive, Settings.TRAIN_SET_SIZE)
            logger.info(state.addr, f"🚂 {len(box)} ballots -> train set of {len(state.train_set)}: {state.train_set}")
        except EarlyStopException:
            return None
        return StageFactory.get_stage("TrainStage" if state.addr in state.train_set else "WaitAggregatedModelsStage")
