#!/usr/bin/env python
"""Headline benchmark: FL rounds/sec of an 8-node federation on N B200s (BASELINE.json, config 2).

    python bench.py --gpus N --steps K --warmup W            (N > 1: launched by torch.distributed.run, one rank per GPU)

One *step* = one full federated round of the whole federation (vote -> evaluate -> local epoch -> partial-model
gossip -> FedAvg -> diffusion) through the public API: ``Node(...)``, ``connect``, ``set_start_learning(rounds, 1)``.
The federation is fixed (8 nodes, CNN on CIFAR-10-shaped synthetic data, fully connected, FedAvg, every node in the
train set) and is spread over the N GPUs ("strong" scaling; N = 8 is exactly "one node per B200").
N = 1 uses the in-memory protocol in one process; N > 1 uses the unix-socket protocol between ranks for the control
plane (B2_BENCH_PROTOCOL=grpc selects gRPC) and CUDA-IPC slab descriptors + peer loads over NVLink for the data plane.

`value` is device-timed (CUDA events on the launching stream, barrier + synchronize on both sides, max over ranks)
with device-resident data; `e2e` repeats the same K rounds with every node's partition coming from pinned host
memory each round (H2D) and each round's losses/metrics read back (D2H), timed by the host clock.  The H2D copies are
double-buffered and issued one call ahead on a copy stream (TorchLearner._tensors): one copy of every partition per round, inside the
timed region, overlapping the previous call's compute and the round's control plane.
"""

from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

if os.environ.get("B2_SWITCH_INTERVAL"):  # experiment: interpreter-lock hand-off latency between node threads (default 5 ms)
    sys.setswitchinterval(float(os.environ["B2_SWITCH_INTERVAL"]))

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

TOTAL_NODES = 8
SAMPLES_PER_NODE = 6250  # CIFAR-10's 50 000 training images split IID over the 8 nodes
TEST_PER_NODE = 1250  # its 10 000 test images, likewise
BATCH = 64
METRIC = "fl_rounds_per_sec"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--model", default="cnn", choices=["cnn", "mlp", "resnet18"])
    ap.add_argument("--nodes", type=int, default=TOTAL_NODES)
    ap.add_argument("--nodes-per-gpu", type=int, default=0, help="weak scaling: this many nodes on every GPU (overrides --nodes)")
    ap.add_argument("--dtype", default="tf32", choices=["tf32", "bf16"], help="tf32 = the reference's precision (fp32 storage, TF32 convolutions, fp32-class linear layers)")
    ap.add_argument("--topology", default="line", choices=["line", "full"], help="line = what the reference arm can run (its own test topology); full = BASELINE config 2 as worded")
    ap.add_argument("--also-full", action="store_true", help="after the timed line-topology rounds, complete the mesh and time the same number of rounds again (reported under 'full_mesh')")
    ap.add_argument("--config", type=int, default=2, choices=[2, 3, 4], help="BASELINE.json config: 2 = CNN / FedAvg / 8 nodes (the headline), 3 = ResNet-18 / 32 nodes / random topology / FedMedian + Krum under attack, 4 = ResNet-18 / Scaffold + FedProx / 64 nodes with drop + rejoin")
    ap.add_argument("--no-coop", action="store_true", help="keep the reference's pairwise partial-model gossip even when the whole federation trains (default: one fused collective per round)")
    ap.add_argument("--no-e2e", action="store_true")
    a = ap.parse_args()
    if a.nodes_per_gpu > 0:
        a.nodes = a.nodes_per_gpu * max(1, int(os.environ.get("WORLD_SIZE", a.gpus)))
    return a


DTYPE_NAMES = {"tf32": "fp32 (tf32 conv, fp32 linear)", "bf16": "bf16"}
MODEL_NAMES = {"cnn": "CNN (3 conv + 2 fc, 0.55M params) on CIFAR-10-shaped data", "mlp": "MLP 784-256-128-10 on MNIST-shaped data", "resnet18": "ResNet-18 CIFAR"}


def common_config(args, n_nodes: int) -> dict:
    """Config keys both arms print with identical values when the runs are the same experiment."""
    return {
        "model": MODEL_NAMES[args.model], "nodes": n_nodes, "train_set_size": n_nodes, "topology": args.topology, "aggregator": "FedAvg",
        "samples_per_node": SAMPLES_PER_NODE, "global_batch": BATCH * n_nodes, "batch_per_node": BATCH, "local_epochs": 1, "seq_len": None,
        "optimizer": "Adam lr 1e-3, fresh per fit", "step": "one federated round of all nodes",
    }


# ------------------------------------------------------------------------------------------------ clocks
class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md recipe)."""

    Q = "index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self) -> None:
        self.proc = None
        self.lines = []

    def start(self) -> None:
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self) -> None:
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self, gpus: int) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            try:
                if int(f[0]) >= gpus:
                    continue
                sm.append(float(f[1]))
                mx.append(float(f[2]))
                for name, v in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"], f[4:8]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            except Exception:
                continue
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------------------ our arm
def make_model(name: str, device, seed: int, dtype: str = "tf32"):
    from myfyp_b200.models import cnn_model, mlp_model, resnet18_model

    if name == "cnn":
        return cnn_model(device=device, seed=seed, precision=dtype)
    return {"mlp": mlp_model, "resnet18": resnet18_model}[name](device=device, seed=seed)


def make_data(name: str, n_nodes: int):
    from myfyp_b200.learning.dataset import P2PFLDataset, RandomIIDPartitionStrategy

    shape = (28, 28) if name == "mlp" else (3, 32, 32)
    data = P2PFLDataset.synthetic(shape, 10, SAMPLES_PER_NODE * n_nodes, TEST_PER_NODE * n_nodes, seed=0)
    return data.generate_partitions(n_nodes, RandomIIDPartitionStrategy)


def run_ours(args) -> None:
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    multi = world > 1
    local_rank = local_rank % torch.cuda.device_count()  # several ranks may share a GPU when debugging on a small box
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if os.environ.get("B2_BENCH_DEBUG"):
        import faulthandler

        os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
        fh = open(os.path.join(ROOT, "gpurun_out", f"bench_hang_rank{rank}.txt"), "w")
        faulthandler.dump_traceback_later(int(os.environ["B2_BENCH_DEBUG"]), file=fh, exit=True)
    shared_gpu = world > torch.cuda.device_count()  # debugging only: NCCL refuses two ranks on one GPU
    if multi:
        if shared_gpu:
            dist.init_process_group("gloo")
        else:
            dist.init_process_group("cpu:gloo,cuda:nccl", device_id=dev)
    red_dev = torch.device("cpu") if shared_gpu else dev

    from myfyp_b200.communication.protocols.grpc.grpc_communication_protocol import GrpcCommunicationProtocol
    from myfyp_b200.communication.protocols.memory.memory_communication_protocol import InMemoryCommunicationProtocol
    from myfyp_b200.learning.aggregators import FedAvg
    from myfyp_b200.node import Node
    from myfyp_b200.ops import native
    from myfyp_b200.parallel.fabric import Fabric
    from myfyp_b200.settings import Settings
    from myfyp_b200.utils.utils import set_fast_settings

    native.load()
    n_nodes = args.nodes
    set_fast_settings(train_set_size=n_nodes)
    Settings.USE_SSL = False
    Settings.TTL = max(Settings.TTL, n_nodes + 2)  # a flood must cross the whole line (reference: examples/mnist.py:153-156)
    Settings.AMOUNT_LAST_MESSAGES_SAVED = max(Settings.AMOUNT_LAST_MESSAGES_SAVED, 200 * n_nodes)
    if os.environ.get("B2_BENCH_DEBUG"):
        from myfyp_b200.management.logger import logger

        Settings.LOG_DIR = os.path.join(ROOT, "gpurun_out", f"logs_rank{rank}")
        logger.enable_file_logging()
        logger.set_level("DEBUG")
    torch.manual_seed(1234 + rank)

    owner = [i * world // n_nodes for i in range(n_nodes)]  # node i lives on rank owner[i]
    mine = [i for i in range(n_nodes) if owner[i] == rank]
    parts = make_data(args.model, n_nodes)
    if multi:
        job = os.environ.get("MASTER_PORT", "29500")
        if os.environ.get("B2_BENCH_PROTOCOL", "uds") == "grpc":
            addrs = [f"unix:///tmp/b2bench-{job}-{i}.sock" for i in range(n_nodes)]
            for i in mine:
                try:
                    os.unlink(addrs[i][len("unix://"):])
                except OSError:
                    pass
            proto = GrpcCommunicationProtocol
        else:  # the box-local control plane: pickled frames over unix sockets
            from myfyp_b200.communication.protocols.uds.uds_communication_protocol import UdsCommunicationProtocol

            addrs = [f"uds:///tmp/b2bench-{job}-{i}.sock" for i in range(n_nodes)]
            proto = UdsCommunicationProtocol
    else:
        addrs = [f"node-{i}" for i in range(n_nodes)]
        proto = InMemoryCommunicationProtocol

    models = {i: make_model(args.model, dev, seed=i, dtype=args.dtype) for i in mine}
    if multi:
        from myfyp_b200.parallel.ipc import ArenaPool, register_addresses

        slot = max(m.layout.total for m in models.values()) if models else 64
        arena = ArenaPool(rank, world, dev, slot_elems=slot, slots=max(4 * len(mine) * 3, 8))
        arena.exchange()
        Fabric.get().attach_arena(arena)
        register_addresses({addrs[i]: owner[i] for i in range(n_nodes)})

    nodes = {i: Node(models[i], parts[i], address=addrs[i], aggregator=FedAvg(), protocol=proto, batch_size=BATCH, lr=1e-3) for i in mine}
    for n in nodes.values():
        n.start()
    if multi:
        dist.barrier()
    hub = None
    if not args.no_coop:
        from myfyp_b200.parallel.coop import CoopHub

        total_elems = next(iter(models.values())).layout.total if models else 64
        hub = CoopHub.configure(rank, world, dev, total_elems, [addrs[i] for i in mine], addrs)

    def wire(topology: str) -> None:
        """line: i - i+1 (the reference's own test topology, test/node_test.py:92-96); full: i -> j for i < j.  The handshake makes
        every link symmetric; the heartbeat flood then tells everybody about everybody (non-direct neighbours)."""
        for i in mine:
            for j in (range(i + 1, n_nodes) if topology == "full" else range(i + 1, min(i + 2, n_nodes))):
                nodes[i].connect(addrs[j])
        deadline = time.time() + 60
        want_direct = (lambda i: n_nodes - 1) if topology == "full" else (lambda i: 1 if i in (0, n_nodes - 1) else 2)
        while any(len(n.get_neighbors(only_direct=True)) < min(want_direct(i), n_nodes - 1) or len(n.get_neighbors(only_direct=False)) < n_nodes - 1 for i, n in nodes.items()):
            if time.time() > deadline:
                raise RuntimeError("topology did not converge")
            time.sleep(0.02)
        if multi:
            dist.barrier()

    wire(args.topology)

    flush_buf = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2

    def flush_l2() -> None:
        flush_buf.zero_()

    first = nodes[mine[0]] if mine else None
    if first is not None:
        first.learner.pre_fit_hook = flush_l2  # one L2 flush per round per GPU

    def barrier() -> None:
        torch.cuda.synchronize(dev)
        if multi:
            dist.barrier()

    def run_rounds(rounds: int) -> None:
        hist = {i: len(n.learning_workflow.history) for i, n in nodes.items()}
        if rank == 0:
            nodes[0].set_start_learning(rounds=rounds, epochs=1)
        t_end = time.time() + 600
        for i, n in nodes.items():
            while not (len(n.learning_workflow.history) > hist[i] and n.learning_workflow.finished):
                if time.time() > t_end:
                    raise RuntimeError(f"node {i} did not finish: {n.learning_workflow.history[-3:]}")
                time.sleep(0.001)

    def set_host_mode(flag: bool) -> None:
        for n in nodes.values():
            n.learner.data_on_host = flag
            n.learner.h2d_bytes = n.learner.d2h_bytes = 0

    def allmax(x: float) -> float:
        if not multi:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=red_dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def allsum(x: float) -> float:
        if not multi:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=red_dev)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    # ---- warm-up (W rounds, untimed), then EXACTLY K timed rounds, device-timed
    barrier()
    run_rounds(args.warmup)
    barrier()
    sampler = ClockSampler()
    if rank == 0:
        sampler.start()
    native.reset_launches()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    for n in nodes.values():
        n.learning_workflow.timings.clear()
    from myfyp_b200.management.logger import logger as _lg

    _lg.reset_metrics()
    from myfyp_b200.management import trace as _trace

    _trace.clear()
    e0.record(torch.cuda.current_stream(dev))
    run_rounds(args.steps)
    e1.record(torch.cuda.current_stream(dev))
    if _trace.ENABLED:  # B2_TRACE=1: control-plane timeline of the timed rounds, one file per rank
        os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
        _trace.dump(os.path.join(ROOT, "gpurun_out", f"trace_rank{rank}.json"))
    barrier()
    stage_ms = {}
    if nodes:
        from collections import defaultdict

        acc_t = defaultdict(float)
        for n in nodes.values():
            for name, dt in n.learning_workflow.timings:
                acc_t[name] += dt
        stage_ms = {k: round(1e3 * v / len(nodes) / args.steps, 2) for k, v in acc_t.items()}
        stage_ms.update({k: round(sum(v) / max(len(v), 1), 2) for k, v in _lg.get_device_times().items()})
        if os.environ.get("B2_BENCH_SERIES"):  # per-round series of one key, e.g. cohort/fit_graph_on_device
            stage_ms["series"] = [round(v, 2) for v in _lg.get_device_times().get(os.environ["B2_BENCH_SERIES"], [])]
        # control-plane messages handled by this rank's nodes, per round (all nodes of the rank together)
        stage_ms.update({k: round(v / args.steps, 1) for k, v in _lg.get_bytes().items() if k.startswith("msgs/")})
    ms = allmax(e0.elapsed_time(e1))
    launches = int(allsum(native.launches()))
    clocks = sampler.stop(world) if rank == 0 else {}

    # ---- end-to-end: inputs from pinned host memory every round, results read back every round
    e2e = None
    if not args.no_e2e:
        set_host_mode(True)
        barrier()
        run_rounds(max(1, min(args.warmup, 2)))
        set_host_mode(True)  # reset byte counters
        barrier()
        t0 = time.perf_counter()
        run_rounds(args.steps)
        barrier()
        wall = allmax(time.perf_counter() - t0)
        h2d = allsum(sum(n.learner.h2d_bytes for n in nodes.values())) / args.steps
        d2h = allsum(sum(n.learner.d2h_bytes for n in nodes.values())) / args.steps
        e2e = {"value": args.steps / wall, "unit": "rounds/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h), "ms_per_step": 1e3 * wall / args.steps,
               "inputs": "pinned host memory -> device every round, double-buffered, copy issued one call ahead on a copy stream"}
        set_host_mode(False)

    # ---- optional second measurement: the same federation as a full mesh (BASELINE config 2 as worded)
    full_mesh = None
    if args.also_full and args.topology != "full":
        try:
            wire("full")
            barrier()
            run_rounds(max(2, args.warmup))
            barrier()
            f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            f0.record(torch.cuda.current_stream(dev))
            run_rounds(args.steps)
            f1.record(torch.cuda.current_stream(dev))
            barrier()
            fms = allmax(f0.elapsed_time(f1))
            full_mesh = {"value": args.steps / (fms / 1e3), "unit": "rounds/s", "ms_per_step": fms / args.steps, "topology": "full"}
        except Exception as e:  # noqa: BLE001 - the headline line must still be printed
            full_mesh = {"error": f"{type(e).__name__}: {e}"[:200]}

    stale = int(allsum(arena.stale_refused)) if multi else 0
    cohort = None
    if args.model == "cnn" and args.dtype == "tf32":
        from myfyp_b200.models.grouped_cnn import CohortTrainer

        trs = list(CohortTrainer._instances.values())
        if trs:
            cohort = {"requests": sum(t.batched for t in trs), "launches": sum(t.batches for t in trs)}

    for n in nodes.values():
        n.stop()
    if multi:
        dist.barrier()
    if rank == 0:
        value = args.steps / (ms / 1e3)
        base = None
        try:
            with open(os.path.join(ROOT, "BASELINE.json")) as f:
                base = json.load(f).get("published", {}).get(METRIC)
        except Exception:
            pass
        out = {
            "metric": METRIC, "value": value, "unit": "rounds/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
            "higher_is_better": True, "scaling": "weak" if args.nodes_per_gpu > 0 else "strong", "vs_baseline": (value / base) if base else None,
            "dtype": DTYPE_NAMES[args.dtype] if args.model == "cnn" else ("bf16" if args.model == "mlp" else "bf16 operands, fp32 accumulate"),
            "data": "synthetic", "impl": "ours",
            "config": {
                **common_config(args, n_nodes),
                "precision": "fp32 storage; convolutions on tcgen05 kind::tf32 (operands rounded to tf32); linear layers 3xTF32 (fp32-class); loss / optimizer fp32" if args.dtype == "tf32" and args.model == "cnn" else "bf16 operands, fp32 accumulation, fp32 master weights",
                "nodes_per_gpu": n_nodes / world, "protocol": (os.environ.get("B2_BENCH_PROTOCOL", "uds") + "+cuda-ipc") if multi else "in-memory",
                "parallelism": f"fed-dp{n_nodes}x{world}gpu", "l2": "flushed (256 MiB write) once per round per GPU",
                "aggregation": ("cooperative: one K-way kernel per GPU" + (" + in-switch (NVLS multimem) sum over the GPUs" if multi and hub is not None and hub.comm is not None else "")) if hub is not None and hub.src is not None else "pairwise partial-model gossip + per-node K-way pull kernel",
            },
            "clocks": clocks, "gpu_launches": launches, "stage_ms_per_round_rank0": stage_ms,
        }
        if full_mesh is not None:
            out["full_mesh"] = full_mesh
        if cohort is not None:
            out["cohort_rank0"] = cohort
        if multi:
            out["stale_descriptors_refused"] = stale
        if hub is not None:
            out["coop_rank0"] = {"rounds_cooperative": hub.successes, "fallbacks": hub.fallbacks}
        if e2e is not None:
            out["e2e"] = e2e
        print(json.dumps(out), flush=True)
    if multi:
        dist.destroy_process_group()


def run_scenarios(args) -> None:
    """Configs 3 and 4: one process drives every GPU of the job (in-memory control plane, peer access between the GPUs), the
    other ranks of a torchrun launch only keep the barrier.  A step = one federated round; `steps` rounds are timed after the
    first one (which holds the graph captures), `value` = the slowest scenario of the config."""
    import torch

    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("gloo")
    try:
        if rank == 0:
            from benchmarks.configs import run
            from myfyp_b200.learning.aggregators import FedMedian, FedProx, Krum, Scaffold

            rounds = args.steps + 1
            n = args.nodes if args.nodes != TOTAL_NODES else (32 if args.config == 3 else 64)
            sampler = ClockSampler()
            sampler.start()
            res = []
            if args.config == 3:
                ts, byz = min(8, n), max(1, n // 8)
                for nm, agg in (("c3-fedmedian", FedMedian), ("c3-krum", lambda: Krum(f=2, m=ts - 4))):
                    res.append(run(nm, n, "resnet18", agg, "random", rounds, 256, ts, "cuda", byzantine=byz, max_gpus=args.gpus))
            else:
                faults = {"drop": [(n - 1, 1), (n - 2, 1)], "rejoin": [(n - 1, 3)]}
                for nm, agg, kw in (("c4-scaffold", lambda: Scaffold(global_lr=1.0, variate_scaling="paper"), {"optimizer": "sgd", "lr": 0.1}), ("c4-fedprox", lambda: FedProx(mu=0.01), None)):
                    res.append(run(nm, n, "resnet18", agg, "random", max(rounds, 6), 256, min(8, n), "cuda", faults=faults, max_gpus=args.gpus, learner_kw=kw))
            clocks = sampler.stop(args.gpus)
            value = min(r["rounds_per_s"] for r in res)
            print(json.dumps({
                "metric": METRIC, "value": value, "unit": "rounds/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": 1, "ms_per_step": 1e3 / value, "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": "bf16 operands, fp32 accumulate (ResNet-18 native step)", "data": "synthetic", "impl": "ours",
                "config": {"baseline_config": args.config, "model": MODEL_NAMES["resnet18"], "nodes": n, "topology": "random", "train_set_size": min(8, n), "protocol": "in-memory, one process, all GPUs",
                           "timing": "wall clock from the end of round 1 to the end of the last round (many streams on many GPUs)"},
                "clocks": clocks, "gpu_launches": sum(r["kernel_launches"] for r in res), "scenarios": res,
            }), flush=True)
    finally:
        if world > 1:
            import torch.distributed as dist

            dist.barrier()
            dist.destroy_process_group()


def main() -> None:
    args = parse()
    if args.config != 2 and args.impl == "ours":
        run_scenarios(args)
        return
    if args.impl == "reference":
        from benchmarks.reference_arm import run_reference

        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
