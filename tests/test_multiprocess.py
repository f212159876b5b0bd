"""Two OS processes, one node each, real gRPC over unix sockets (no in-process shortcut), pickled weights on the wire."""

import multiprocessing as mp
import os
import tempfile
import time

import numpy as np


def _worker(rank, addrs, q, rounds, proto_name="grpc"):
    import torch

    torch.set_num_threads(1)
    from myfyp_b200.communication.protocols.grpc.grpc_communication_protocol import GrpcCommunicationProtocol
    from myfyp_b200.learning.dataset import P2PFLDataset, RandomIIDPartitionStrategy
    from myfyp_b200.models import mlp_model
    from myfyp_b200.node import Node
    from myfyp_b200.utils.utils import set_test_settings, wait_convergence

    set_test_settings()
    from myfyp_b200.management.logger import logger

    logger.set_level("ERROR")
    data = P2PFLDataset.synthetic_mnist(400, 100, seed=0).generate_partitions(2, RandomIIDPartitionStrategy)[rank]
    from myfyp_b200.communication.protocols.uds.uds_communication_protocol import UdsCommunicationProtocol

    proto = GrpcCommunicationProtocol if proto_name == "grpc" else UdsCommunicationProtocol
    node = Node(mlp_model(seed=rank), data, address=addrs[rank], protocol=proto, batch_size=32)
    node.start()
    q.put(("up", rank))
    if rank == 0:
        t_conn = time.time() + 60  # the peer process may still be importing torch
        while not node.connect(addrs[1]):
            assert time.time() < t_conn, "could not reach the peer process"
            time.sleep(0.5)
    wait_convergence([node], 1, wait=10, only_direct=True)
    if rank == 0:
        node.set_start_learning(rounds=rounds, epochs=1)
    t_end = time.time() + 90
    while not (len(node.learning_workflow.history) > 1 and node.learning_workflow.finished):
        assert time.time() < t_end, node.learning_workflow.history
        time.sleep(0.02)
    params = node.get_model().get_parameters()
    q.put(("done", rank, [float(np.sum(p)) for p in params], node.learning_workflow.history))
    time.sleep(1.0)
    node.stop()


import pytest


@pytest.mark.parametrize("proto_name", ["grpc", "uds"])
def test_two_processes_over_sockets(proto_name):
    ctx = mp.get_context("spawn")
    d = tempfile.mkdtemp()
    addrs = [f"unix://{d}/n{i}.sock" if proto_name == "grpc" else f"uds://{d}/n{i}.sock" for i in range(2)]
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, addrs, q, 2, proto_name)) for r in range(2)]
    for p in procs:
        p.start()
    results = {}
    deadline = time.time() + 150
    while len(results) < 2 and time.time() < deadline:
        try:
            msg = q.get(timeout=1)
        except Exception:
            if any(p.exitcode not in (None, 0) for p in procs):
                break
            continue
        if msg[0] == "done":
            results[msg[1]] = msg[2:]
    for p in procs:
        p.join(20)
        if p.is_alive():
            p.terminate()
    assert len(results) == 2, "a worker died"
    (s0, h0), (s1, h1) = results[0], results[1]
    np.testing.assert_allclose(s0, s1, rtol=1e-4, atol=1e-3)  # same aggregated model on both sides
    assert h0.count("RoundFinishedStage") == 2 and h1.count("RoundFinishedStage") == 2


def _rank_worker(rank, addrs, q, rounds, per_rank, topology="full"):
    """`per_rank` nodes in this process (local shortcut between them), unix-socket frames to the other rank - the shape of
    `bench.py` with several nodes per GPU process."""
    import torch

    torch.set_num_threads(1)
    from myfyp_b200.communication.protocols.uds.uds_communication_protocol import UdsCommunicationProtocol
    from myfyp_b200.learning.dataset import P2PFLDataset, RandomIIDPartitionStrategy
    from myfyp_b200.management.logger import logger
    from myfyp_b200.models import mlp_model
    from myfyp_b200.node import Node
    from myfyp_b200.settings import Settings
    from myfyp_b200.utils.utils import set_test_settings

    set_test_settings()
    Settings.TRAIN_SET_SIZE = len(addrs)
    logger.set_level("ERROR")
    parts = P2PFLDataset.synthetic_mnist(100 * len(addrs), 25 * len(addrs), seed=0).generate_partitions(len(addrs), RandomIIDPartitionStrategy)
    mine = list(range(rank * per_rank, (rank + 1) * per_rank))
    nodes = {i: Node(mlp_model(seed=i), parts[i], address=addrs[i], protocol=UdsCommunicationProtocol, batch_size=32) for i in mine}
    for n in nodes.values():
        n.start()
    q.put(("up", rank))
    t_conn = time.time() + 90
    for i, n in nodes.items():  # full mesh: every node connects to every node with a larger index; line: to the next one only
        for j in range(i + 1, len(addrs) if topology == "full" else min(i + 2, len(addrs))):
            while not n.connect(addrs[j]):
                assert time.time() < t_conn, "could not reach a peer"
                time.sleep(0.3)
    t_end = time.time() + 60
    while any(len(n.get_neighbors(only_direct=False)) < len(addrs) - 1 for n in nodes.values()):
        assert time.time() < t_end, "neighbour tables did not converge"
        time.sleep(0.05)
    q.put(("ready", rank))
    if rank == 0:
        time.sleep(1.0)  # let the other rank finish its own connects
        nodes[0].set_start_learning(rounds=rounds, epochs=1)
    t_end = time.time() + 120
    while not all(len(n.learning_workflow.history) > 1 and n.learning_workflow.finished for n in nodes.values()):
        assert time.time() < t_end, {i: n.learning_workflow.history[-3:] for i, n in nodes.items()}
        time.sleep(0.02)
    for i, n in nodes.items():
        q.put(("done", i, [float(np.sum(p)) for p in n.get_model().get_parameters()], n.learning_workflow.history.count("RoundFinishedStage")))
    time.sleep(1.0)
    for n in nodes.values():
        n.stop()


@pytest.mark.parametrize("topology,per_rank", [("full", 2), ("line", 3)])
def test_two_ranks_with_several_nodes_each_over_unix_sockets(topology, per_rank):
    """``line``: every flooded message crosses in-process hops (forwarded on the delivering thread, one shared "seen" set) AND a
    socket hop (the set leaves the process as the wire's "seen" list) on its way from one end to the other."""
    ctx = mp.get_context("spawn")
    d = tempfile.mkdtemp()
    rounds = 3
    addrs = [f"uds://{d}/n{i}.sock" for i in range(2 * per_rank)]
    q = ctx.Queue()
    procs = [ctx.Process(target=_rank_worker, args=(r, addrs, q, rounds, per_rank, topology)) for r in range(2)]
    for p in procs:
        p.start()
    results = {}
    deadline = time.time() + 240
    while len(results) < len(addrs) and time.time() < deadline:
        try:
            msg = q.get(timeout=1)
        except Exception:
            if any(p.exitcode not in (None, 0) for p in procs):
                break
            continue
        if msg[0] == "done":
            results[msg[1]] = msg[2:]
    for p in procs:
        p.join(20)
        if p.is_alive():
            p.terminate()
    assert len(results) == len(addrs), "a worker died"
    for i in range(1, len(addrs)):
        np.testing.assert_allclose(results[0][0], results[i][0], rtol=1e-4, atol=1e-3)  # one aggregated model everywhere
        assert results[i][1] == rounds
