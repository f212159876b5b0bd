#!/usr/bin/env python
"""BASELINE.json configs 1-4 as runnable scenarios (single process, every visible GPU, in-memory control plane).

    python benchmarks/configs.py --config 3 [--nodes 32] [--rounds 3] [--model resnet18] [--samples-per-node 512]

  1  MLP, MNIST-shaped, 2 nodes, in-memory protocol, FedAvg, CPU (plumbing)
  2  CNN, CIFAR-10-shaped, 8 nodes, full mesh, FedAvg                      (bench.py is the timed, multi-process version)
  3  ResNet-18, 32 nodes, random topology, FedMedian and Krum (with sign-flip attackers to show why they exist)
  4  ResNet-18, Scaffold and FedProx, 64 nodes, nodes dropped mid-run and rejoining

Nodes are placed round-robin on the visible GPUs (many per GPU) by the fabric; aggregation kernels pull peers' slabs
over NVLink (peer access).  Prints one JSON line per scenario: rounds/s (wall clock over the timed rounds, first round
excluded - it contains the CUDA-graph captures), final accuracy of honest nodes, kernel launches, fault events.
"""

from __future__ import annotations

import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def run(name, n_nodes, model_name, aggregator_factory, topology, rounds, samples, train_set, device_mode, byzantine=0, faults=None, seed=0, batch=64, graphs=None, max_gpus=None, learner_kw=None):
    import numpy as np
    import torch

    from myfyp_b200.communication.protocols.memory.memory_communication_protocol import InMemoryCommunicationProtocol
    from myfyp_b200.learning.dataset import P2PFLDataset, RandomIIDPartitionStrategy
    from myfyp_b200.management.logger import logger
    from myfyp_b200.models import cnn_model, mlp_model, resnet18_model
    from myfyp_b200.node import Node
    from myfyp_b200.ops import native
    from myfyp_b200.parallel.fabric import Fabric
    from myfyp_b200.settings import Settings
    from myfyp_b200.utils.fault_injector import FaultInjector, make_byzantine
    from myfyp_b200.utils.topologies import TopologyFactory
    from myfyp_b200.utils.utils import set_fast_settings

    set_fast_settings(train_set_size=train_set)
    Settings.TTL = max(10, n_nodes)
    Settings.AMOUNT_LAST_MESSAGES_SAVED = 100000
    Settings.VOTE_TIMEOUT = 120
    Settings.AGGREGATION_TIMEOUT = 120
    Settings.HEARTBEAT_TIMEOUT = 120
    if faults:  # a real failure detector: evict within a few seconds, do not sit out minute-long time-outs
        Settings.HEARTBEAT_PERIOD, Settings.HEARTBEAT_TIMEOUT = 0.5, 3
        Settings.VOTE_TIMEOUT, Settings.AGGREGATION_TIMEOUT = 15, 30
    logger.reset_metrics()
    shape = (28, 28) if model_name == "mlp" else (3, 32, 32)
    data = P2PFLDataset.synthetic(shape, 10, samples * n_nodes, max(64, samples // 4) * n_nodes, seed=seed)
    parts = data.generate_partitions(n_nodes, RandomIIDPartitionStrategy, seed=seed)
    mk = {"mlp": mlp_model, "cnn": cnn_model, "resnet18": resnet18_model}[model_name]
    gpus = torch.cuda.device_count() if device_mode == "cuda" else 0
    if max_gpus:
        gpus = min(gpus, max_gpus)
    if graphs is None:  # all three models train through their hand-written steps: one CUDA graph per local epoch
        graphs = True
    Settings.GOSSIP_EXIT_ON_X_EQUAL_ROUNDS = 60
    if gpus > 1:
        Fabric.get().enable_peer_access()
    nodes = []
    for i in range(n_nodes):
        dev = f"cuda:{i % gpus}" if gpus else "cpu"
        nodes.append(Node(mk(device=dev, seed=seed + i), parts[i], address=f"{name}-{i}", aggregator=aggregator_factory(), protocol=InMemoryCommunicationProtocol, batch_size=batch, use_cuda_graph=graphs, **(learner_kw or {})))
    for n in nodes:
        n.start()
    for n in nodes[n_nodes - byzantine :]:
        make_byzantine(n, "sign_flip", scale=5.0)
    TopologyFactory.connect_nodes(TopologyFactory.generate_matrix(topology, n_nodes, p=0.15, seed=seed), nodes)
    deadline = time.time() + 120
    while any(len(n.get_neighbors(only_direct=False)) < n_nodes - 1 for n in nodes):
        if time.time() > deadline:
            raise RuntimeError("topology did not converge")
        time.sleep(0.05)
    fi = FaultInjector()
    late = []
    if faults:
        for j, (idx, rnd) in enumerate(faults.get("drop", [])):
            # the first victim dies wherever it is in that round, the others only if they are training at that moment
            fi.drop_at(nodes[idx], round=rnd, stage=None if j == 0 else "TrainStage")
        for idx, at_round in faults.get("rejoin", []):
            def make(idx=idx):
                n = Node(mk(device=f"cuda:{idx % gpus}" if gpus else "cpu", seed=1000 + idx), parts[idx], address=f"{name}-re-{idx}", aggregator=aggregator_factory(), protocol=InMemoryCommunicationProtocol, batch_size=batch, use_cuda_graph=graphs, **(learner_kw or {}))
                late.append(n)
                return n
            targets = [nodes[j].addr for j in range(min(4, n_nodes)) if j != idx]
            fi.rejoin_at(make, targets, nodes[0], at_round)
    native.reset_launches()
    t0 = time.perf_counter()
    nodes[0].set_start_learning(rounds=rounds, epochs=1)
    # time from the end of round 1 (graph captures happen in it) to the end of the last round
    t_r1 = None
    dropped = {idx for idx, _ in (faults or {}).get("drop", [])}
    alive = [n for i, n in enumerate(nodes) if i not in dropped]
    t_end = time.time() + 3600
    while not all(n.learning_workflow.finished for n in alive):
        if t_r1 is None and all((n.state.round or 0) >= 1 or n.learning_workflow.finished for n in alive):
            t_r1 = time.perf_counter()
        if time.time() > t_end:
            raise RuntimeError("scenario did not finish")
        time.sleep(0.002)
    if gpus:
        for g in range(gpus):
            torch.cuda.synchronize(g)
    t1 = time.perf_counter()
    glob = logger.get_global_logs().get(Settings.EXPERIMENT_NAME, {})
    honest = [n for i, n in enumerate(nodes) if i < n_nodes - byzantine and i not in dropped]
    # final accuracy: every honest node evaluates the model it ended up with (the per-round metric of a node is only logged in
    # rounds in which it was voted into the train set, so "last logged value" can be many rounds old)
    acc = []
    for n in honest:
        try:
            acc.append(float(n.learner.evaluate().get("test_metric")))
        except Exception:  # noqa: BLE001
            pass
    acc = [a for a in acc if a == a]
    out = {
        "scenario": name, "nodes": n_nodes, "gpus": gpus, "model": model_name, "topology": topology, "train_set_size": train_set, "rounds": rounds,
        "samples_per_node": samples, "byzantine": byzantine, "rounds_per_s": round((rounds - 1) / (t1 - t_r1), 3) if t_r1 and rounds > 1 else round(rounds / (t1 - t0), 3),
        "total_s": round(t1 - t0, 2), "honest_acc_mean": round(float(np.mean(acc)), 4) if acc else None, "honest_acc_min": round(float(np.min(acc)), 4) if acc else None,
        "kernel_launches": native.launches(), "fault_events": [(e[0], e[1], e[2]) for e in fi.events],
        "rejoined_rounds": {n.addr: n.learning_workflow.history.count("RoundFinishedStage") for n in late},
    }
    # wall time per stage, mean over the nodes that finished, per round (ms): where a round goes at this scale
    acc_t: dict = {}
    for n in alive:
        for st_name, dt in n.learning_workflow.timings:
            acc_t[st_name] = acc_t.get(st_name, 0.0) + dt
    out["stage_ms_per_round"] = {k: round(1e3 * v / max(len(alive), 1) / rounds, 1) for k, v in acc_t.items()}
    print(json.dumps(out), flush=True)
    for n in nodes + late:
        try:
            n.stop()
        except Exception:  # noqa: BLE001
            pass
    time.sleep(0.2)
    return out


def main() -> None:
    import torch

    from myfyp_b200.learning.aggregators import FedAvg, FedMedian, FedProx, Krum, Scaffold

    ap = argparse.ArgumentParser()
    ap.add_argument("--config", type=int, required=True, choices=[1, 2, 3, 4])
    ap.add_argument("--nodes", type=int, default=None)
    ap.add_argument("--rounds", type=int, default=None)
    ap.add_argument("--model", default=None)
    ap.add_argument("--samples-per-node", type=int, default=None)
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    mode = "cuda" if torch.cuda.is_available() and a.config != 1 else "cpu"
    if a.rounds is None:
        a.rounds = 6 if a.config == 4 else 3
    res = []
    if a.config == 1:
        res.append(run("c1-mlp-fedavg", a.nodes or 2, a.model or "mlp", FedAvg, "line", a.rounds, a.samples_per_node or 600, 4, "cpu"))
    elif a.config == 2:
        n = a.nodes or 8
        res.append(run("c2-cnn-fedavg", n, a.model or "cnn", FedAvg, "full", a.rounds, a.samples_per_node or 6250, n, mode))
    elif a.config == 3:
        n = a.nodes or 32
        byz = max(1, n // 8)
        ts = min(8, n)
        for nm, agg in (("c3-fedmedian", FedMedian), ("c3-krum", lambda: Krum(f=2, m=ts - 4)), ("c3-fedavg-under-attack", FedAvg)):
            res.append(run(nm, n, a.model or "resnet18", agg, "random", a.rounds, a.samples_per_node or 512, ts, mode, byzantine=byz))
    else:
        n = a.nodes or 64
        faults = {"drop": [(n - 1, 1), (n - 2, 1)], "rejoin": [(n - 1, 3)]}  # two nodes die in round 1, one comes back in round 3
        # Scaffold's control variates estimate gradients from (x - y) / (K lr): that is SGD arithmetic (Karimireddy et al., Alg. 1), so
        # its clients run SGD; FedProx keeps the reference's Adam
        for nm, agg, kw in (("c4-scaffold", lambda: Scaffold(global_lr=1.0, variate_scaling="paper"), {"optimizer": "sgd", "lr": 0.1}), ("c4-fedprox", lambda: FedProx(mu=0.01), None)):
            res.append(run(nm, n, a.model or "resnet18", agg, "random", a.rounds, a.samples_per_node or 256, min(8, n), mode, faults=faults, learner_kw=kw))
    if a.out:
        os.makedirs(os.path.dirname(os.path.abspath(a.out)) or ".", exist_ok=True)
        with open(a.out, "w") as f:
            json.dump(res, f, indent=1)


if __name__ == "__main__":
    main()
