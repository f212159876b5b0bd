"""Logger, metric stores, node monitor and the web-services client (the reference ships `test/management/p2pfl_web_test.py`
as a docstring-only stub; its store shapes are pinned by `node_test.py:128` through `logger.get_global_logs()`)."""

import datetime
import http.server
import json
import threading
import time

import pytest

from myfyp_b200.experiment import Experiment
from myfyp_b200.management.logger import logger
from myfyp_b200.management.metric_storage import GlobalMetricStorage, LocalMetricStorage
from myfyp_b200.management.node_monitor import NodeMonitor, gpu_metrics
from myfyp_b200.management.web_services import P2pflWebServices, P2pflWebServicesError
from myfyp_b200.settings import Settings


def test_metric_store_shapes():
    loc, glo = LocalMetricStorage(), GlobalMetricStorage()
    loc.add_log("exp", 0, "loss", "n1", 1.0, 0)
    loc.add_log("exp", 0, "loss", "n1", 0.5, 1)
    loc.add_log("exp", 1, "loss", "n2", 0.4, 0)
    assert loc.get_all_logs() == {"exp": {0: {"n1": {"loss": [(0, 1.0), (1, 0.5)]}}, 1: {"n2": {"loss": [(0, 0.4)]}}}}
    assert loc.get_experiment_round_node_logs("exp", 0, "n1") == {"loss": [(0, 1.0), (1, 0.5)]}
    glo.add_log("exp", 0, "acc", "n1", 0.1)
    glo.add_log("exp", 1, "acc", "n1", 0.6)
    glo.add_log("exp", 1, "acc", "n1", 0.7)  # one value per round: a second log of the same round does not add a point
    series = glo.get_experiment_node_logs("exp", "n1")["acc"]
    assert [r for r, _ in series] == [0, 1]
    loc.clear()
    glo.clear()
    assert loc.get_all_logs() == {} and glo.get_all_logs() == {}


def test_logger_routes_metrics_by_step_and_drops_unregistered_nodes():
    logger.reset_metrics()
    logger.register_node("mgmt-a")
    try:
        logger.log_metric("mgmt-a", "m", 1.0)  # no experiment yet: dropped silently (reference logger.py:286-308)
        assert logger.get_global_logs() == {}
        logger.experiment_started("mgmt-a", Experiment("exp-m", 3))
        logger.log_metric("mgmt-a", "test_metric", 0.5)  # global (per round)
        logger.log_metric("mgmt-a", "train_loss", 2.0, step=0)  # local (per step)
        logger.log_metric("nobody", "x", 1.0)  # unknown node: dropped
        assert logger.get_global_logs()["exp-m"]["mgmt-a"]["test_metric"] == [(0, 0.5)]
        assert logger.get_local_logs()["exp-m"][0]["mgmt-a"]["train_loss"] == [(0, 2.0)]
        logger.log_device_time("k", 1.5)
        logger.log_bytes("msgs/x", 2)
        assert logger.get_device_times()["k"] == [1.5] and logger.get_bytes()["msgs/x"] == 2
        lvl = logger.get_level()
        logger.set_level("ERROR")
        assert logger.get_level_name(logger.get_level()) == "ERROR"
        logger.set_level(lvl)
    finally:
        logger.unregister_node("mgmt-a")
        logger.reset_metrics()


def test_node_monitor_reports_cpu_ram_net():
    keep = Settings.RESOURCE_MONITOR_PERIOD
    Settings.RESOURCE_MONITOR_PERIOD = 0.05
    seen = {}
    mon = NodeMonitor("mgmt-b", lambda node, k, v, t: seen.setdefault(k, []).append((node, v, t)))
    try:
        mon.start()
        deadline = time.time() + 5
        while time.time() < deadline and not {"cpu", "ram", "net_in", "net_out"} <= set(seen):
            time.sleep(0.02)
        assert {"cpu", "ram", "net_in", "net_out"} <= set(seen)
        node, v, t = seen["ram"][0]
        assert node == "mgmt-b" and 0 <= v <= 100 and isinstance(t, datetime.datetime)
        assert isinstance(gpu_metrics(), dict)  # {} without a GPU / NVML
    finally:
        mon.stop()
        Settings.RESOURCE_MONITOR_PERIOD = keep


class _Handler(http.server.BaseHTTPRequestHandler):
    calls = []

    def do_POST(self):  # noqa: N802
        body = json.loads(self.rfile.read(int(self.headers["Content-Length"])) or b"{}")
        _Handler.calls.append((self.path, self.headers.get("x-api-key"), body))
        if self.path == "/fail":
            self.send_response(500)
            self.end_headers()
            self.wfile.write(b"boom")
            return
        self.send_response(201 if self.path == "/node" else 200)
        self.send_header("Content-Type", "application/json")
        self.end_headers()
        self.wfile.write(json.dumps({"node_id": 42}).encode())

    def log_message(self, *a):
        pass


def test_web_services_client_against_a_local_stub():
    pytest.importorskip("requests")
    srv = http.server.HTTPServer(("127.0.0.1", 0), _Handler)
    threading.Thread(target=srv.serve_forever, daemon=True).start()
    try:
        web = P2pflWebServices(f"http://127.0.0.1:{srv.server_port}", "secret")
        now = datetime.datetime.now()
        web.send_log(now, "n", 20, "dropped: node not registered yet")
        assert _Handler.calls == []
        web.register_node("n", is_simulated=False)
        assert web.node_id["n"] == 42
        web.send_log(now, "n", 20, "hello")
        web.send_local_metric("exp", 1, "loss", "n", 0.5, 3)
        web.send_global_metric("exp", 1, "acc", "n", 0.9)
        web.send_system_metric("n", "cpu", 12.0, now)
        paths = [c[0] for c in _Handler.calls]
        assert paths == ["/node", "/node-log", "/node-metric/local", "/node-metric/global", "/node-metric/system"]
        assert all(c[1] == "secret" for c in _Handler.calls)
        assert _Handler.calls[2][2] == {"node_id": 42, "exp": "exp", "round": 1, "metric_name": "loss", "metric_value": 0.5, "step": 3}
        with pytest.raises(P2pflWebServicesError):
            web._post("fail", {})
        with pytest.raises(NotImplementedError):
            web.get_pending_actions()
        web.unregister_node("n")
        assert "n" not in web.node_id
    finally:
        srv.shutdown()


def _rank_logs(rank, port, q):
    import os

    import torch.distributed as dist

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE="2")
    dist.init_process_group("gloo", rank=rank, world_size=2)
    from myfyp_b200.management.decorators import RayP2PFLogger
    from myfyp_b200.management.logger import logger

    addr = f"node-of-rank-{rank}"
    logger.register_node(addr)
    logger.experiment_started(addr, __import__("myfyp_b200.experiment", fromlist=["Experiment"]).Experiment("experiment", 3))
    logger.log_metric(addr, "test_metric", 0.5 + rank)             # global store
    logger.log_metric(addr, "train_loss", 1.0 - 0.1 * rank, step=7)  # local store
    view = RayP2PFLogger(logger)
    g, l = view.get_global_logs(), view.get_local_logs()
    q.put((rank, sorted(g["experiment"].keys()), sorted(l["experiment"][0].keys()), sorted(view.get_global_logs(all_ranks=False)["experiment"].keys())))
    dist.destroy_process_group()


def test_cross_rank_logger_view_merges_the_ranks_metric_stores():
    """The role of the reference's Ray logger actor (ray_logger.py:32-250): every process sees the metrics of all nodes."""
    import multiprocessing as mp
    import socket

    from myfyp_b200.management.decorators import _merge_logs

    assert _merge_logs({"e": {"a": {"m": [(0, 1.0)]}}}, {"e": {"a": {"m": [(0, 1.0), (1, 2.0)]}, "b": {"m": [(0, 3.0)]}}}) == {"e": {"a": {"m": [(0, 1.0), (1, 2.0)]}, "b": {"m": [(0, 3.0)]}}}
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    ps = [ctx.Process(target=_rank_logs, args=(r, port, q)) for r in range(2)]
    [p.start() for p in ps]
    got = sorted(q.get(timeout=120) for _ in range(2))
    [p.join(30) for p in ps]
    both = ["node-of-rank-0", "node-of-rank-1"]
    for rank, glob, loc, mine in got:
        assert glob == both and loc == both, got   # every rank sees both nodes
        assert mine == [f"node-of-rank-{rank}"]    # the un-gathered view is the local one
