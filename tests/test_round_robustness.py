"""Unit tests of the round bookkeeping that keeps a federation in step when nodes run at different speeds or die:
per-round ballots, ballots counted whoever cast them, early model offers, lingering announcements, the small seams
(framework registry, singleton, trace, cross-GPU ordering helper on CPU)."""

import threading
import time

import numpy as np
import pytest
import torch

from myfyp_b200.communication.commands.message.vote_train_set_command import VoteTrainSetCommand
from myfyp_b200.communication.commands.weights.partial_model_command import PartialModelCommand
from myfyp_b200.communication.protocols.client import Client
from myfyp_b200.communication.protocols.gossiper import Gossiper
from myfyp_b200.learning.aggregators import FedAvg
from myfyp_b200.models import mlp_model
from myfyp_b200.node_state import NodeState


@pytest.fixture(autouse=True)
def _restore_settings():
    from myfyp_b200.settings import Settings

    keep = (Settings.TRAIN_SET_SIZE, Settings.VOTE_TIMEOUT)
    yield
    Settings.TRAIN_SET_SIZE, Settings.VOTE_TIMEOUT = keep


def _state(round_=3):
    st = NodeState("me")
    st.set_experiment("experiment", 10, round_)
    return st


def test_next_round_ballots_do_not_overwrite_current_ones():
    """Reference keeps one dict (vote_train_set_command.py:60-71): a fast peer's ballot for round r+1 replaces its ballot for
    round r on a node that is still tallying r, which then elects a different train set than everybody else."""
    st = _state(3)
    cmd = VoteTrainSetCommand(st)
    cmd.execute("a", 3, "x", "10", "y", "5")
    cmd.execute("a", 4, "z", "99")  # a is already one round ahead
    cmd.execute("b", 3, "y", "7")
    assert st.train_set_votes == {"a": {"x": 10, "y": 5}, "b": {"y": 7}}
    assert st.train_set_votes_next == {"a": {"z": 99}}
    cmd.execute("c", 9, "q", "1")  # far future: ignored
    assert "c" not in st.train_set_votes and "c" not in st.train_set_votes_next
    st.increase_round()
    assert st.round == 4 and st.train_set_votes == {"a": {"z": 99}} and st.train_set_votes_next == {}


class _Proto:
    """The slice of a protocol the vote stage and the partial-model command use."""

    def __init__(self, neighbors):
        self.neighbors = list(neighbors)
        self.sent = []

    def get_neighbors(self, only_direct=False):
        return {n: None for n in self.neighbors}

    def build_msg(self, cmd, args=None, round=None):
        return {"cmd": cmd, "args": list(args or []), "round": round}

    def broadcast(self, msg, node_list=None):
        self.sent.append(msg)

    def broadcast_async(self, msg, coalesce=False, linger=0.0):
        self.sent.append((msg, coalesce, linger))


def test_ballot_of_a_node_that_left_the_neighbour_table_still_counts():
    """Whether a dead node's ballot is in the tally must not depend on when a node noticed the death."""
    from myfyp_b200.settings import Settings
    from myfyp_b200.stages.base_node.vote_train_set_stage import VoteTrainSetStage

    Settings.TRAIN_SET_SIZE = 2
    Settings.VOTE_TIMEOUT = 5
    results = []
    for neighbours in (["a", "b", "dead"], ["a", "b"]):  # one node still lists the dead peer, the other has evicted it
        st = _state(1)
        # the dead node's ballot decides the election (our own random ballot weighs at most 1000 per candidate)
        st.train_set_votes = {"a": {"a": 10**6}, "b": {"a": 1}, "dead": {"b": 10**7}}
        proto = _Proto(neighbours)
        nxt = VoteTrainSetStage.execute(state=st, communication_protocol=proto)
        assert nxt is not None
        results.append(list(st.train_set))
    assert results[0] == results[1] == ["b", "a"]


def test_early_model_offers_are_kept_and_filed_when_training_starts():
    st = _state(2)
    agg = FedAvg("me")
    model = mlp_model(seed=0)

    class _Learner:
        def get_model(self):
            return model

    proto = _Proto(["a", "b"])
    cmd = PartialModelCommand(st, lambda: None, agg, proto, _Learner())
    blob = model.encode_parameters()
    # (1) the sender is a whole round ahead, (2) same round but we have not started collecting yet
    cmd.execute("a", 3, weights=blob, contributors=["a"], num_samples=10)
    assert agg.get_aggregated_models() == [] and len(st.early_models) == 1
    st.increase_round()
    st.train_set = ["me", "a", "b"]
    cmd.execute("b", 3, weights=blob, contributors=["b"], num_samples=10)
    assert len(st.early_models) == 2 and agg.get_aggregated_models() == []
    agg.set_nodes_to_aggregate(st.train_set)
    cmd.absorb_early()
    assert sorted(agg.get_aggregated_models()) == ["a", "b"] and st.early_models == []
    # the sender evidently holds what it sent
    assert st.models_aggregated["a"] == ["a"] and st.models_aggregated["b"] == ["b"]
    # announcements go out asynchronously, coalesced
    assert proto.sent and all(isinstance(m, tuple) and m[1] is True for m in proto.sent)
    # a stale offer (older round) is dropped, not stashed
    cmd.execute("a", 1, weights=blob, contributors=["a"], num_samples=10)
    assert st.early_models == []


class _RecordingClient(Client):
    def __init__(self):
        self.sent = []
        self.lock = threading.Lock()

    def build_message(self, cmd, args=None, round=None):
        return {"cmd": cmd, "args": args or [], "round": round}

    def build_weights(self, cmd, round, serialized_model, contributors=None, weight=1):
        return {}

    def send(self, nei, msg, create_connection=False, raise_error=False, remove_on_error=True):
        with self.lock:
            self.sent.append((time.monotonic(), nei, msg["args"]))

    def broadcast(self, msg, node_list=None):
        pass


def test_gossiper_linger_collapses_a_burst_and_a_final_message_is_not_delayed():
    client = _RecordingClient()
    g = Gossiper("me", client, period=0.0, messages_per_period=1000)
    g.start()
    try:
        t0 = time.monotonic()
        for i in range(5):  # a burst of coalescable announcements, held back 0.15 s
            g.add_message({"cmd": "models_aggregated", "args": [str(i)], "round": 1}, ["a", "b"], coalesce_key=("models_aggregated", 1), linger=0.15)
            time.sleep(0.01)
        time.sleep(0.05)
        assert client.sent == []  # still lingering
        time.sleep(0.3)
        assert [(n, a) for _, n, a in client.sent] == [("a", ["4"]), ("b", ["4"])]  # one message per neighbour, the newest
        assert client.sent[0][0] - t0 >= 0.14
        client.sent.clear()
        # a message that must not wait (linger 0) supersedes a lingering one and goes out at once
        g.add_message({"cmd": "models_aggregated", "args": ["x"], "round": 2}, ["a"], coalesce_key=("models_aggregated", 2), linger=5.0)
        g.add_message({"cmd": "models_aggregated", "args": ["final"], "round": 2}, ["a"], coalesce_key=("models_aggregated", 2), linger=0.0)
        time.sleep(0.2)
        assert [(n, a) for _, n, a in client.sent] == [("a", ["final"])]
        # plain messages are never delayed
        client.sent.clear()
        g.add_message({"cmd": "vote", "args": ["v"], "round": 2}, ["b"])
        time.sleep(0.1)
        assert [(n, a) for _, n, a in client.sent] == [("b", ["v"])]
    finally:
        g.stop()


def test_small_seams():
    from myfyp_b200 import ops
    from myfyp_b200.learning.frameworks import Framework, learner_for, register_learner
    from myfyp_b200.management import trace as tr
    from myfyp_b200.utils.singleton import SingletonMeta

    assert learner_for(Framework.PYTORCH.value).__name__ == "TorchLearner"
    try:
        learner_for("no-such-framework")
        raise AssertionError("expected ValueError")
    except ValueError as e:
        assert "register_learner" in str(e)
    register_learner("mine", lambda: dict)
    assert learner_for("mine") is dict

    class One(metaclass=SingletonMeta):
        pass

    assert One() is One()

    from myfyp_b200.learning.simulation import try_init_learner_with_ray, try_init_learner_with_simulation
    from myfyp_b200.management.decorators import AsyncLogger, FileLogger, LoggerDecorator, RayP2PFLogger, WebP2PFLogger  # noqa: F401
    from myfyp_b200.utils.check_ray import ray_installed

    assert try_init_learner_with_ray is try_init_learner_with_simulation and ray_installed() is False

    class _L:
        calls = []

        def enable_async(self):
            self.calls.append("async")

        def enable_file_logging(self, d=None):
            self.calls.append(("file", d))

        def info(self, node, msg):
            self.calls.append((node, msg))

    stacked = FileLogger(AsyncLogger(_L()), "/tmp/x")
    stacked.info("n", "hello")
    assert _L.calls == ["async", ("file", "/tmp/x"), ("n", "hello")]

    ops.order_after_producers([torch.zeros(3)], torch.device("cpu"))  # CPU: nothing to order, must not touch CUDA
    was = tr.ENABLED
    try:
        tr.ENABLED = True
        tr.clear()
        tr.trace("n", "e")
        import json
        import os
        import tempfile

        path = os.path.join(tempfile.mkdtemp(), "t.json")
        tr.dump(path)
        ev = json.load(open(path))
        assert len(ev) == 1 and ev[0][1:] == ["n", "e"]
    finally:
        tr.ENABLED = was
        tr.clear()


def test_aggregates_are_computed_where_the_node_lives():
    """`Aggregator.local_device`: the model this node contributed decides the device (CPU here; on a multi-GPU box the GPU)."""
    base = mlp_model(seed=0)
    ref = base.get_parameters()
    agg = FedAvg("me")
    ms = [base.build_copy(params=[p + i for p in ref], num_samples=1, contributors=[c]) for i, c in enumerate(["other", "me"])]
    assert agg.local_device(ms) == ms[1].slab.device
    out = agg.aggregate(ms)
    np.testing.assert_allclose(out.get_parameters()[0], ref[0] + 0.5, atol=1e-6)
    assert FedAvg("nobody").local_device(ms) == ms[0].slab.device


def test_a_peer_that_lacks_our_model_is_offered_exactly_that_model():
    """TrainStage's offer policy: own model alone first (it can never overlap what the peer holds), other nodes' models only to
    peers that already have ours and only after `FORWARD_GRACE` (the origin usually delivers directly meanwhile)."""
    from myfyp_b200.settings import Settings
    from myfyp_b200.stages.base_node.train_stage import TrainStage

    keep = Settings.FORWARD_GRACE
    Settings.FORWARD_GRACE = 0.05
    st = _state(1)
    st.train_set = ["me", "a", "b"]
    base = mlp_model(seed=0)
    ref = base.get_parameters()
    agg = FedAvg("me")
    agg.set_nodes_to_aggregate(st.train_set)
    agg.add_model(base.build_copy(params=ref, num_samples=10, contributors=["me"]))
    agg.add_model(base.build_copy(params=[p + 1 for p in ref], num_samples=10, contributors=["a"]))
    captured = {}

    class _P(_Proto):
        def gossip_weights(self, early_stopping_fn, get_candidates_fn, status_fn, model_fn, **kw):
            captured.update(candidates=get_candidates_fn, model=model_fn)

        def build_weights(self, cmd, round, model, contributors=None, weight=1):
            return {"cmd": cmd, "round": round, "contributors": list(contributors), "weight": weight}

        def fabric_reachable(self, nei):
            return False

    try:
        TrainStage._TrainStage__gossip_model_aggregation(st, _P(["a", "b"]), agg)
        assert sorted(captured["candidates"]()) == ["a", "b"]
        first = captured["model"]("b")  # b has announced nothing: it gets our own model, not the {me, a} aggregate
        assert first["contributors"] == ["me"] and first["cmd"] == "partial_model"
        assert captured["model"]("b") is None  # the same offer is not repeated right away
        st.models_aggregated["b"] = ["me"]  # b has ours now and still lacks a's
        assert captured["model"]("b") is None  # pure forwarding waits for the grace period
        time.sleep(0.08)
        fwd = captured["model"]("b")
        assert fwd is not None and fwd["contributors"] == ["a"]
        st.models_aggregated["b"] = ["me", "a"]
        assert "b" not in captured["candidates"]() or captured["model"]("b") is None
    finally:
        Settings.FORWARD_GRACE = keep


def test_coalescing_never_replaces_a_forwarded_message_of_another_node():
    """The coalescing tag stays inside the sender's queue.  It used to travel with the message: a forwarder's own next
    announcement then replaced the foreign announcement still queued for forwarding, and in multi-hop topologies nodes
    waited for announcements that had been overwritten on the way (found with tools/stall_hunt.py on a 12-node ring)."""
    client = _RecordingClient()
    sent_msgs = []
    orig = client.send

    def send(nei, msg, **kw):
        sent_msgs.append((nei, dict(msg)))
        return orig(nei, msg, **kw)

    client.send = send
    g = Gossiper("me", client, period=0.0, messages_per_period=1000)
    # a foreign announcement queued for forwarding, as it arrives off the wire (no tag), then our own coalescable one
    g.add_message({"cmd": "models_aggregated", "args": ["theirs"], "round": 1, "source": "other"}, ["b"])
    g.add_message({"cmd": "models_aggregated", "args": ["ours-1"], "round": 1, "source": "me"}, ["b"], coalesce_key=("me", "models_aggregated", 1), linger=0.05)
    g.add_message({"cmd": "models_aggregated", "args": ["ours-2"], "round": 1, "source": "me"}, ["b"], coalesce_key=("me", "models_aggregated", 1), linger=0.05)
    g.start()
    try:
        time.sleep(0.3)
        assert [m["args"] for _, m in sent_msgs] == [["theirs"], ["ours-2"]]
        assert all("_ckey" not in m for _, m in sent_msgs)
    finally:
        g.stop()


def test_model_gossip_stall_exit_counts_periods_not_wake_ups():
    """GOSSIP_EXIT_ON_X_EQUAL_ROUNDS is about time (the reference's rounds are one period apart): a storm of unrelated state
    notifications must not end the model gossip early, but x periods without progress must."""
    from myfyp_b200.settings import Settings

    class _Client:
        def send(self, nei, msg, create_connection=False):
            pass

    old = Settings.GOSSIP_EXIT_ON_X_EQUAL_ROUNDS
    Settings.GOSSIP_EXIT_ON_X_EQUAL_ROUNDS = 4
    try:
        g = Gossiper("me", _Client(), period=0.0, messages_per_period=1000)
        wakes = []

        def wait_fn(left):  # every wait is cut short at once, as a busy NodeState.notify() does
            wakes.append(time.monotonic())
            time.sleep(0.002)

        t0 = time.monotonic()
        g.gossip_weights(lambda: False, lambda: ["peer"], lambda: "no progress", lambda nei: None, period=0.1, create_connection=False, wait_fn=wait_fn)
        took = time.monotonic() - t0
        assert len(wakes) > 40, "the loop was woken many times"
        assert 0.38 <= took < 1.5, f"exit after 4 periods of 0.1 s without progress, not after 4 wake-ups (took {took:.3f} s)"
    finally:
        Settings.GOSSIP_EXIT_ON_X_EQUAL_ROUNDS = old
