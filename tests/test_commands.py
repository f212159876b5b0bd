"""What each wire command does to a node (the reference ships `test/communication/commands_test.py` docstring-only).
Wire names are part of the protocol: they are asserted literally."""

import numpy as np
import pytest

from myfyp_b200.communication.commands.message.heartbeat_command import HeartbeatCommand
from myfyp_b200.communication.commands.message.metrics_command import MetricsCommand
from myfyp_b200.communication.commands.message.model_initialized_command import ModelInitializedCommand
from myfyp_b200.communication.commands.message.models_aggregated_command import ModelsAggregatedCommand
from myfyp_b200.communication.commands.message.models_ready_command import ModelsReadyCommand
from myfyp_b200.communication.commands.message.start_learning_command import StartLearningCommand
from myfyp_b200.communication.commands.message.stop_learning_command import StopLearningCommand
from myfyp_b200.communication.commands.message.vote_train_set_command import VoteTrainSetCommand
from myfyp_b200.communication.commands.weights.full_model_command import FullModelCommand
from myfyp_b200.communication.commands.weights.init_model_command import InitModelCommand
from myfyp_b200.communication.commands.weights.partial_model_command import PartialModelCommand
from myfyp_b200.experiment import Experiment
from myfyp_b200.learning.aggregators import FedAvg
from myfyp_b200.management.logger import logger
from myfyp_b200.models import mlp_model
from myfyp_b200.node_state import NodeState


def _state(round_=2):
    st = NodeState("me")
    st.set_experiment("experiment", 10, round_)
    return st


def test_wire_names():
    names = [c.get_name() for c in (HeartbeatCommand, StartLearningCommand, StopLearningCommand, ModelInitializedCommand, VoteTrainSetCommand,
                                    ModelsAggregatedCommand, ModelsReadyCommand, MetricsCommand, InitModelCommand, PartialModelCommand, FullModelCommand)]
    assert names == ["beat", "start_learning", "stop_learning", "model_initialized", "vote_train_set", "models_aggregated", "models_ready",
                     "metrics", "init_model", "partial_model", "add_model"]


def test_heartbeat_and_start_learning_forward_their_arguments():
    beats, starts = [], []

    class _HB:
        def beat(self, source, time):
            beats.append((source, time))

    HeartbeatCommand(_HB()).execute("a", -1, "123.5")
    assert beats == [("a", 123.5)]
    with pytest.raises(ValueError):
        HeartbeatCommand(_HB()).execute("a", -1)
    cmd = StartLearningCommand(lambda rounds, epochs, **kw: starts.append((rounds, epochs, kw)))
    cmd.execute("a", -1, "5", "2")
    cmd.execute("a", -1, "5", "2", "3")  # rejoin invitation: the experiment is already in round 3
    assert starts == [(5, 2, {}), (5, 2, {"start_round": 3})]
    with pytest.raises(ValueError):
        cmd.execute("a", -1, "5")


def test_round_bookkeeping_commands():
    st = _state(2)
    ModelInitializedCommand(st).execute("a", 0)
    assert st.nei_status["a"] == -1
    ready = ModelsReadyCommand(st)
    ready.execute("a", 2)
    ready.execute("b", 1)  # b announces the PREVIOUS round's model: it is filed as round 1, i.e. b still needs round 2's model
    ready.execute("c", 0)  # long past: ignored
    ready.execute("a", 1)  # an older announcement overtaken by a newer one does not move a backwards
    ready.execute("d", 3)  # a peer that is ahead of us
    assert st.nei_status == {"a": 2, "b": 1, "d": 3}
    agg = ModelsAggregatedCommand(st)
    agg.execute("a", 2, "x", "y")
    agg.execute("a", 2, "y", "z")  # merged, not overwritten: announcements can overtake each other
    agg.execute("b", 3, "q")  # the sender is a round ahead: kept for when we get there
    agg.execute("c", 9, "w")  # ignored
    assert st.models_aggregated == {"a": ["x", "y", "z"]} and st.models_aggregated_next == {"b": ["q"]}
    st.increase_round()
    assert st.models_aggregated == {"b": ["q"]} and st.models_aggregated_next == {}
    idle = NodeState("idle")
    ModelsReadyCommand(idle).execute("a", 0)
    VoteTrainSetCommand(idle).execute("a", 0, "x", "1")  # not learning: ignored
    assert idle.nei_status == {} and idle.train_set_votes == {}


def test_metrics_command_logs_the_peers_metrics_as_global_metrics():
    logger.reset_metrics()
    logger.register_node("peer-m")
    try:
        logger.experiment_started("peer-m", Experiment("exp-c", 3))
        MetricsCommand(_state()).execute("peer-m", 0, "test_loss", "0.25", "test_metric", "0.9")
        got = logger.get_global_logs()["exp-c"]["peer-m"]
        assert got["test_loss"] == [(0, 0.25)] and got["test_metric"] == [(0, 0.9)]
    finally:
        logger.unregister_node("peer-m")
        logger.reset_metrics()


class _Learner:
    def __init__(self, model):
        self.model, self.interrupted, self.set_calls = model, False, 0

    def get_model(self):
        return self.model

    def set_model(self, m):
        self.set_calls += 1
        self.model.set_parameters(m)

    def interrupt_fit(self):
        self.interrupted = True


def test_stop_learning_clears_everything():
    st = _state(2)
    st.train_set = ["me", "a"]
    agg = FedAvg("me")
    agg.set_nodes_to_aggregate(["me", "a"])
    learner = _Learner(mlp_model(seed=0))
    StopLearningCommand(st, agg, learner).execute("a", 2)
    assert learner.interrupted and st.round is None and st.train_set == [] and agg.get_missing_models() == set()


def test_init_and_full_model_commands_adopt_once():
    model = mlp_model(seed=0)
    other = mlp_model(seed=1)
    blob = other.encode_parameters()
    stopped = []
    # init_model: only while the node still waits for its initial model (lock held)
    st = _state(0)
    learner = _Learner(model)
    init = InitModelCommand(st, lambda: stopped.append(1), FedAvg("me"), learner, None)
    assert st.model_initialized_lock.locked()
    init.execute("a", 0, weights=blob)
    assert learner.set_calls == 1 and not st.model_initialized_lock.locked()
    init.execute("a", 0, weights=blob)  # already initialised: ignored
    assert learner.set_calls == 1
    np.testing.assert_allclose(model.get_parameters()[0], other.get_parameters()[0])
    # add_model (full aggregated model): only while waiting for it, once per round
    st2 = _state(1)
    learner2 = _Learner(mlp_model(seed=2))
    full = FullModelCommand(st2, lambda: stopped.append(1), FedAvg("me"), learner2, None)
    full.execute("a", 1, weights=blob)  # event is set (not waiting): not expected
    assert learner2.set_calls == 0
    st2.aggregated_model_event.clear()
    full.execute("a", 7, weights=blob)  # wrong round
    assert learner2.set_calls == 0
    full.execute("a", 1, weights=blob)
    full.execute("b", 1, weights=blob)  # a second neighbour pushes the same aggregate: ignored
    assert learner2.set_calls == 1 and st2.aggregated_model_event.is_set()
    with pytest.raises(ValueError):
        full.execute("a", 1)
    assert stopped == []
    # an aggregate pushed before the node waits for it (still tallying votes, or a round behind) is kept and adopted on entry
    st3 = _state(1)
    learner3 = _Learner(mlp_model(seed=3))
    early = FullModelCommand(st3, lambda: stopped.append(1), FedAvg("me"), learner3, None)
    early.execute("a", 1, weights=blob)  # same round, not waiting yet
    assert learner3.set_calls == 0 and st3.early_full_model[0] == 1
    st3.aggregated_model_event.clear()  # WaitAggregatedModelsStage starts waiting ...
    assert early.adopt_early() and learner3.set_calls == 1 and st3.early_full_model is None
    early.execute("a", 2, weights=blob)  # the sender is a whole round ahead
    assert st3.early_full_model[0] == 2 and learner3.set_calls == 1
    assert early.adopt_early() is False  # ... not this round's model: dropped, nothing adopted
    assert learner3.set_calls == 1
