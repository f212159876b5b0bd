"""Learner, callbacks, simulation pool, CLI and settings behaviour on the CPU path."""

import subprocess
import sys

import numpy as np
import pytest
import torch

from myfyp_b200.learning.aggregators import FedProx, Scaffold
from myfyp_b200.learning.callbacks import CallbackFactory, FedProxCallback, SCAFFOLDCallback
from myfyp_b200.learning.dataset import P2PFLDataset
from myfyp_b200.learning.learner import LearnerFactory
from myfyp_b200.learning.simulation import FitPool, VirtualNodeLearner
from myfyp_b200.learning.torch_learner import TorchLearner
from myfyp_b200.models import cnn_model, mlp_model
from myfyp_b200.settings import Settings


def test_fit_and_evaluate_learn_and_log_metrics():
    from myfyp_b200.experiment import Experiment
    from myfyp_b200.management.logger import logger

    data = P2PFLDataset.synthetic_mnist(512, 128, seed=3)
    model = mlp_model(seed=0)
    assert LearnerFactory.create_learner(model) is TorchLearner
    learner = TorchLearner(model, data, "learner-node", batch_size=32, extra_metrics=True)
    logger.register_node("learner-node", False)
    logger.experiment_started("learner-node", Experiment("experiment", 1))
    before = learner.evaluate()
    learner.set_epochs(2)
    learner.fit()
    after = learner.evaluate()
    assert after["test_metric"] > max(0.9, before["test_metric"]) and after["test_loss"] < before["test_loss"]
    assert {"test_precision", "test_recall", "test_f1"} <= set(after) and after["test_f1"] > 0.9
    assert model.get_contributors() == ["learner-node"] and model.get_num_samples() == 512
    local = logger.get_local_logs()["experiment"][0]["learner-node"]["train_loss"]
    assert len(local) == 2 * 16 and local[-1][1] < local[0][1]
    assert "test_metric" in logger.get_global_logs()["experiment"]["learner-node"]


def test_cnn_learner_cpu_path():
    data = P2PFLDataset.synthetic_cifar10(512, 64, seed=1)
    learner = TorchLearner(cnn_model(seed=0), data, "n", batch_size=32, lr=2e-3)
    learner.set_epochs(5)
    before = learner.evaluate()["test_loss"]
    learner.fit()
    after = learner.evaluate()
    assert after["test_loss"] < 0.5 * before and after["test_metric"] > 0.6


def test_interrupt_fit_stops_early():
    data = P2PFLDataset.synthetic_mnist(4096, 64)
    learner = TorchLearner(mlp_model(seed=0), data, "n", batch_size=8)
    learner.set_epochs(50)
    import threading

    t = threading.Thread(target=learner.fit)
    t.start()
    learner.interrupt_fit()
    t.join(20)
    assert not t.is_alive()


def test_fedprox_term_pulls_towards_the_global_model():
    data = P2PFLDataset.synthetic_mnist(256, 32, seed=2)
    drift = {}
    for mu in (0.0, 10.0):
        torch.manual_seed(0)
        model = mlp_model(seed=0)
        start = model.slab.clone()
        learner = TorchLearner(model, data, "n", aggregator=FedProx(mu=mu), batch_size=32, shuffle=False)
        assert [c.get_name() for c in learner.callbacks] == ["fedprox"] and learner.callbacks[0].mu == mu
        learner.set_epochs(3)
        learner.fit()
        drift[mu] = float((model.slab - start).norm())
    assert drift[10.0] < 0.7 * drift[0.0]


def test_scaffold_callback_produces_the_deltas_the_aggregator_needs():
    data = P2PFLDataset.synthetic_mnist(128, 32, seed=2)
    model = mlp_model(seed=0)
    learner = TorchLearner(model, data, "n", aggregator=Scaffold(), batch_size=32, lr=1e-2, optimizer="sgd")
    cb = learner.callbacks[0]
    assert isinstance(cb, SCAFFOLDCallback)
    x = model.slab.clone()
    learner.fit()
    info = model.get_info("scaffold")
    n = model.layout.n_trainable
    torch.testing.assert_close(info["delta_y_i"][:n], (model.slab - x)[:n])
    steps = 4
    torch.testing.assert_close(info["delta_c_i"][:n], ((x - model.slab) / (steps * 1e-2))[:n], rtol=1e-4, atol=1e-5)  # c_i started at 0
    torch.testing.assert_close(cb.c_i, info["delta_c_i"][:n])
    # the wire form is NumPy (reference-compatible pickle)
    import pickle

    blob = pickle.loads(model.encode_parameters())
    wire = blob["additional_info"]["scaffold"]["delta_y_i"]  # the reference's form: one array per state_dict entry
    assert isinstance(wire, list) and [a.shape for a in wire] == [tuple(e.shape) for e in model.layout.in_order()]
    # a global_c coming back from the aggregator is picked up at the next fit
    model.add_info("scaffold", {"global_c": torch.ones_like(model.slab)})
    learner.update_callbacks_with_model_info()
    learner.fit()
    assert float(cb.c.mean()) == 1.0
    with pytest.raises(ValueError):
        CallbackFactory.create_callbacks("pytorch", type("A", (), {"get_required_callbacks": lambda self: ["nope"]})())
    assert FedProxCallback.get_name() == "fedprox"


def test_virtual_learner_admission_control():
    FitPool.configure(1)
    data = P2PFLDataset.synthetic_mnist(64, 16)
    v = VirtualNodeLearner(TorchLearner(mlp_model(seed=0), data, "n"))
    v.set_epochs(1)
    v.fit()
    assert v.evaluate()["test_metric"] >= 0.0 and v.get_framework() == "pytorch"
    v.interrupt_fit()  # the reference raises NotImplementedError here
    FitPool.configure(16)


def test_settings_follow_mutation():
    Settings.HEARTBEAT_TIMEOUT = 50
    assert Settings.WAIT_HEARTBEATS_CONVERGENCE == 10.0  # derived at read time (the reference freezes it at import)
    Settings.WAIT_HEARTBEATS_CONVERGENCE = 0.5
    assert Settings.WAIT_HEARTBEATS_CONVERGENCE == 0.5


def test_cli_lists_and_runs_examples():
    out = subprocess.run([sys.executable, "-m", "myfyp_b200", "experiment", "list"], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0 and "mnist" in out.stdout and "byzantine" in out.stdout
    out = subprocess.run([sys.executable, "-m", "myfyp_b200", "experiment", "run", "nope"], capture_output=True, text=True, timeout=120)
    assert out.returncode == 1
    out = subprocess.run([sys.executable, "-m", "myfyp_b200", "remote"], capture_output=True, text=True, timeout=120)
    assert "Not implemented yet" in out.stdout  # like the reference's login / remote (its launch too; ours works, see below)
    out = subprocess.run(
        [sys.executable, "-m", "myfyp_b200", "experiment", "run", "mnist", "--nodes", "2", "--rounds", "1", "--measure_time", "--log-level", "ERROR", "--samples-per-node", "100"],
        capture_output=True, text=True, timeout=300,
    )
    assert out.returncode == 0 and "seconds" in out.stdout


def test_native_library_loads_without_a_gpu():
    from myfyp_b200.ops import native

    lib = native.load()
    assert lib.b2_version() >= 100
    assert lib.b2_device_count() >= 0


def test_build_entry_point():
    import __graft_entry__ as g

    g.build()


def test_settings_bulk_update_env_and_file(tmp_path, monkeypatch):
    """`Settings.update / update_from_env / load_file` (the reference can only be configured by assigning class attributes)."""
    import json

    from myfyp_b200.settings import Settings

    keep = {k: getattr(Settings, k) for k in ("TRAIN_SET_SIZE", "USE_SSL", "GOSSIP_PERIOD", "LOG_LEVEL")}
    keep_wait = Settings.__dict__.get("_WAIT_HEARTBEATS_CONVERGENCE")
    try:
        Settings.update({"TRAIN_SET_SIZE": "8", "USE_SSL": "false", "GOSSIP_PERIOD": "0.25", "WAIT_HEARTBEATS_CONVERGENCE": 1.5})
        assert (Settings.TRAIN_SET_SIZE, Settings.USE_SSL, Settings.GOSSIP_PERIOD, Settings.WAIT_HEARTBEATS_CONVERGENCE) == (8, False, 0.25, 1.5)
        with pytest.raises(KeyError):
            Settings.update({"TRAIN_SET_SIZ": 3})
        monkeypatch.setenv("MYFYP_B200_TRAIN_SET_SIZE", "5")
        monkeypatch.setenv("MYFYP_B200_NOT_A_SETTING", "1")
        assert Settings.update_from_env() == ["TRAIN_SET_SIZE"] and Settings.TRAIN_SET_SIZE == 5
        path = tmp_path / "s.json"
        path.write_text(json.dumps({"LOG_LEVEL": "ERROR", "TRAIN_SET_SIZE": 2}))
        Settings.load_file(str(path))
        assert Settings.LOG_LEVEL == "ERROR" and Settings.TRAIN_SET_SIZE == 2
        assert "HEARTBEAT_PERIOD" in Settings.names() and "update" not in Settings.names()
    finally:
        for k, v in keep.items():
            setattr(Settings, k, v)
        Settings._WAIT_HEARTBEATS_CONVERGENCE = keep_wait


def test_cli_launch_two_nodes_over_grpc(tmp_path):
    """`python -m myfyp_b200 launch`: one process serves, a second connects, initiates two rounds and exits."""
    import os
    import time

    env = dict(os.environ, MYFYP_B200_LOG_LEVEL="ERROR", MYFYP_B200_HEARTBEAT_PERIOD="0.5")
    a, b = f"unix://{tmp_path}/a.sock", f"unix://{tmp_path}/b.sock"
    server = subprocess.Popen([sys.executable, "-m", "myfyp_b200", "launch", "--address", a, "--samples", "128"], env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    try:
        deadline = time.time() + 240  # a cold `import torch` can take a minute
        while not os.path.exists(f"{tmp_path}/a.sock"):
            assert time.time() < deadline and server.poll() is None, "the serving node did not come up"
            time.sleep(0.2)
        out = subprocess.run([sys.executable, "-m", "myfyp_b200", "launch", "--address", b, "--connect", a, "--samples", "128", "--rounds", "2", "--seed", "1"],
                             env=env, capture_output=True, text=True, timeout=420)
        assert out.returncode == 0, out.stdout + out.stderr
        assert "learning finished" in out.stdout
    finally:
        server.terminate()
        try:
            server.wait(timeout=20)
        except subprocess.TimeoutExpired:
            server.kill()


def test_cohort_trainer_batches_concurrent_requests_and_releases_everybody_on_errors():
    """Host side of the node-batched trainer (models/grouped_cnn.py) without a GPU: requests that arrive within the window run as one
    cohort, `fit` and `evaluate` never share one, a late request becomes the next cohort, and an exception in the launch is raised
    in EVERY waiting learner instead of leaving some of them blocked."""
    import threading
    import time

    from myfyp_b200.models.grouped_cnn import CohortTrainer

    class Fake(CohortTrainer):
        def __init__(self):
            super().__init__(None, None, torch.device("cpu"), 64)
            self.window, self.max_group = 0.25, 4
            self.cohorts = []

        def _execute(self, kind, batch):
            self.cohorts.append((kind, sorted(j.args["who"] for j in batch)))
            if any(j.args.get("boom") for j in batch):
                raise RuntimeError("launch failed")
            time.sleep(0.02)
            for j in batch:
                j.result = (kind, j.args["who"])

    tr = Fake()
    class L:  # stands for a learner (the trainer keeps weak references)
        pass

    members = [L() for _ in range(4)]
    for m in members:
        tr.register(m)
    out, errs = {}, {}

    def call(i, kind, **kw):
        try:
            out[i] = tr.submit(kind, members[i % 4], dict(who=i, **kw))
        except Exception as e:  # noqa: BLE001
            errs[i] = e

    ts = [threading.Thread(target=call, args=(i, "fit")) for i in range(4)]
    for t in ts:
        t.start()
    for t in ts:
        t.join(5)
    assert tr.cohorts == [("fit", [0, 1, 2, 3])] and out == {i: ("fit", i) for i in range(4)}  # all four registered learners asked: one cohort, no waiting out the window

    tr.cohorts.clear()
    out.clear()
    ts = [threading.Thread(target=call, args=(0, "fit")), threading.Thread(target=call, args=(1, "eval")), threading.Thread(target=call, args=(2, "fit"))]
    for t in ts:
        t.start()
    for t in ts:
        t.join(5)
    kinds = sorted(tr.cohorts)
    assert all(len({k}) == 1 for k, _ in kinds) and sorted(w for _, ws in kinds for w in ws) == [0, 1, 2]
    assert not any(k == "fit" and 1 in ws for k, ws in kinds) and not any(k == "eval" and (0 in ws or 2 in ws) for k, ws in kinds)  # kinds never mix

    tr.cohorts.clear()
    out.clear()
    ts = [threading.Thread(target=call, args=(i, "fit"), kwargs=dict(boom=(i == 1))) for i in range(4)]
    for t in ts:
        t.start()
    for t in ts:
        t.join(5)
    assert not any(t.is_alive() for t in ts) and sorted(errs) == [0, 1, 2, 3] and all("launch failed" in str(e) for e in errs.values())
    # the trainer is usable afterwards
    errs.clear()
    call(0, "fit")
    assert out[0] == ("fit", 0) and not errs
