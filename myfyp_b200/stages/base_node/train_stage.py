"""TrainStage (reference: stages/base_node/train_stage.py:44-187).

evaluate -> fit -> add own model -> announce -> gossip partial aggregates inside the train set until every
member covers the train set -> aggregate -> adopt -> announce ``models_ready``.

Data plane: what is gossiped is ``aggregator.get_model(except_nodes=what the peer announced)``
(SURVEY App. C.1).  To a peer on the NVLink fabric the message carries a slab descriptor; the peer's
aggregation kernel pulls it out of our HBM when it reduces.  Off-fabric peers get the reference pickle.
"""

from __future__ import annotations

import time
from typing import Any, List, Set, Type, Union

from myfyp_b200.communication.commands.message.metrics_command import MetricsCommand
from myfyp_b200.communication.commands.message.models_aggregated_command import ModelsAggregatedCommand
from myfyp_b200.communication.commands.message.models_ready_command import ModelsReadyCommand
from myfyp_b200.communication.commands.weights.partial_model_command import PartialModelCommand
from myfyp_b200.learning.aggregators.aggregator import NoModelsToAggregateError
from myfyp_b200.management.logger import logger
from myfyp_b200.management.nvtx import nvtx_range
from myfyp_b200.management.trace import trace
from myfyp_b200.settings import Settings
from myfyp_b200.stages.diffusion import transport_payload as _transport
from myfyp_b200.stages.stage import EarlyStopException, Stage, check_early_stop
from myfyp_b200.stages.stage_factory import StageFactory


class TrainStage(Stage):
    @staticmethod
    def name() -> str:
        return "TrainStage"

    @staticmethod
    def execute(state=None, communication_protocol=None, learner=None, aggregator=None, **kwargs) -> Union[Type[Stage], None]:
        if state is None or communication_protocol is None or aggregator is None or learner is None:
            raise Exception("Invalid parameters on TrainStage.")
        t = [time.perf_counter()]

        def lap(name: str) -> None:  # wall-time split of the stage, read by benchmarks (logger.get_device_times())
            t.append(time.perf_counter())
            logger.log_device_time(f"train_stage/{name}", 1e3 * (t[-1] - t[-2]))
            trace(state.addr, f"lap {name} r{state.round}")

        try:
            check_early_stop(state)
            aggregator.set_nodes_to_aggregate(state.train_set)
            # a member evicted by the failure detector between the vote's validation and this point was reported to an
            # aggregator that did not know the train set yet: check once more, or the round waits for a dead node's model
            alive = set(communication_protocol.get_neighbors(only_direct=False)) | {state.addr}
            aggregator.federation_size = len(alive)
            for dead in [n for n in state.train_set if n not in alive]:
                if aggregator.drop_node(dead):
                    state.train_set = [n for n in state.train_set if n != dead]
                    logger.info(state.addr, f"☠️ {dead} left before training started: aggregating without it.")
            if getattr(state, "absorb_early_models", None) is not None:
                state.absorb_early_models()  # offers from peers that started this round before we did
            check_early_stop(state)
            evaluate_and_broadcast(state, learner, communication_protocol)
            check_early_stop(state)
            lap("evaluate")

            logger.info(state.addr, "🏋️‍♀️ Training...")
            with nvtx_range(f"{state.addr}/fit"):
                learner.fit()
            check_early_stop(state)
            lap("fit")

            if TrainStage.__cooperative_round(state, communication_protocol, learner, aggregator, lap):
                return StageFactory.get_stage("GossipModelStage")

            models_added = aggregator.add_model(learner.get_model().snapshot_model())
            communication_protocol.broadcast(communication_protocol.build_msg(ModelsAggregatedCommand.get_name(), models_added, round=state.round))
            lap("publish")
            TrainStage.__gossip_model_aggregation(state, communication_protocol, aggregator)
            check_early_stop(state)
            lap("gossip")

            with nvtx_range(f"{state.addr}/aggregate"):
                agg_model = aggregator.wait_and_get_aggregation()
            lap("aggregate")
            learner.set_model(agg_model)
            state.aggregate_round = state.round
            communication_protocol.broadcast(communication_protocol.build_msg(ModelsReadyCommand.get_name(), [], round=state.round))
            lap("adopt")
            return StageFactory.get_stage("GossipModelStage")
        except EarlyStopException:
            return None

    @staticmethod
    def __cooperative_round(state, communication_protocol, learner, aggregator, lap) -> bool:
        """Whole federation in the train set, everybody on the fabric: the resident nodes of this GPU aggregate once, together, and
        the GPUs sum their partials in one fused collective (parallel/coop.py) instead of n (n-1) pairwise model offers
        (reference: train_stage.py:120-176 + fedavg.py:61-69).  False = not applicable / did not complete: gossip as usual."""
        from myfyp_b200.parallel.coop import CoopHub

        hub = CoopHub.get()
        if hub is None or not hub.eligible(state.addr, state.train_set, aggregator):
            return False
        if state.round == 0:
            hub.experiment_started(state.addr)
        model = learner.get_model()
        res = hub.aggregate(state.addr, state.round, model.slab, model.get_num_samples())
        if res is None:
            logger.info(state.addr, "Cooperative aggregation did not complete: gossiping this round.")
            return False
        average, total = res
        check_early_stop(state)
        everybody = list(state.train_set)
        state.models_aggregated[state.addr] = everybody
        communication_protocol.broadcast(communication_protocol.build_msg(ModelsAggregatedCommand.get_name(), everybody, round=state.round))
        lap("publish")
        lap("gossip")
        agg_model = model.build_copy(params=average, num_samples=total, contributors=everybody)
        lap("aggregate")
        learner.set_model(agg_model)
        state.aggregate_round = state.round
        communication_protocol.broadcast(communication_protocol.build_msg(ModelsReadyCommand.get_name(), [], round=state.round))
        lap("adopt")
        return True

    @staticmethod
    def __gossip_model_aggregation(state, communication_protocol, aggregator) -> None:
        def early_stopping_fn() -> bool:
            return state.round is None

        def aggregated_by(node: str) -> List[str]:
            return state.models_aggregated.get(node, [])

        def remaining(node: str) -> Set[str]:
            return set(state.train_set) - set(aggregated_by(node))

        def excluded_for(node: str) -> List[str]:
            """Contributors that must not be in an offer to ``node``.  A peer that still lacks OUR OWN model gets exactly that
            model, alone: a single fresh model can never overlap what the peer holds, whereas a partial aggregate built from a
            stale view of the peer (own + others it has meanwhile received directly) is refused as a whole - and our own model
            with it.  (The reference always offers the widest aggregate and relies on its once-per-second retries.)"""
            have = aggregated_by(node)
            if state.addr not in have:
                others = [n for n in state.train_set if n != state.addr]
                if aggregator.peek_offer(others):  # we do hold our own model as a single (we always do after fit)
                    return others
            return have

        def fresh_offer(node: str) -> bool:
            key = (node, frozenset(aggregator.peek_offer(excluded_for(node))))
            return bool(key[1]) and time.monotonic() - offered.get(key, -1e9) >= reoffer_after

        def get_candidates_fn() -> List[str]:
            # the loop ends on the peers' ANNOUNCEMENTS, not on what we offered: an offer can be refused (overlap with
            # something the peer got meanwhile), and only its next announcement tells us what it still lacks
            return [n for n in set(state.train_set) - {state.addr} if len(remaining(n)) != 0]

        def status_fn() -> Any:
            return [(n, aggregated_by(n)) for n in communication_protocol.get_neighbors(only_direct=False) if n in state.train_set]

        # What was already offered to whom in this round.  The loop below is event driven (it runs again as soon as any
        # announcement arrives), so without this memory every wake-up would re-send the same model to every peer that
        # has not finished yet - O(n^3) transfers per round in a full mesh.  An identical offer is repeated only after
        # `reoffer_after` seconds (covers a lost message, like the reference's periodic re-send does).
        offered: dict = {}
        reoffer_after = max(1.0, 20 * Settings.GOSSIP_MODELS_PERIOD)

        arrival: dict = {}  # contributor -> when this loop first saw it in our held set

        def model_fn(node: str) -> Any:
            if not fresh_offer(node):  # nothing new for this peer: do not even run the partial-aggregation kernel
                return None
            excluded = excluded_for(node)
            offer = aggregator.peek_offer(excluded)
            now = time.monotonic()
            for c in offer:
                arrival.setdefault(c, now)
            if state.addr not in offer and any(now - arrival[c] < Settings.FORWARD_GRACE for c in offer):
                return None  # pure forwarding: give the origins a moment to deliver directly (the loop wakes up again)
            try:
                model = aggregator.get_model(excluded)
            except NoModelsToAggregateError:
                logger.debug(state.addr, f"❔ No models to aggregate from {node}.")
                return None
            if state.round is None:
                raise Exception("Round not initialized.")
            offered[(node, frozenset(model.get_contributors()))] = time.monotonic()
            return communication_protocol.build_weights(
                PartialModelCommand.get_name(), state.round, _transport(model, communication_protocol, node), model.get_contributors(), model.get_num_samples()
            )

        communication_protocol.gossip_weights(
            early_stopping_fn, get_candidates_fn, status_fn, model_fn, create_connection=True, wait_fn=state.wait_change
        )


def evaluate_and_broadcast(state, learner, communication_protocol) -> None:
    logger.info(state.addr, "🔬 Evaluating...")
    results = learner.evaluate()
    logger.info(state.addr, f"📈 Evaluated. Results: {results}")
    if len(results) > 0:
        flat = [str(item) for pair in results.items() for item in pair]
        communication_protocol.broadcast(communication_protocol.build_msg(MetricsCommand.get_name(), flat, round=state.round))
