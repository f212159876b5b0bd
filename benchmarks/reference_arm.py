"""Reference arm of bench.py: the UNMODIFIED reference (p2pfl 0.4.0 = PrivEimantas/myFYP) from ``baseline/_ref``.

What runs is the reference's own ``Node`` + stage machine + in-memory protocol + ``FedAvg`` aggregator + ``Settings``
- its public API and stock code path.  None of myfyp_b200 is imported here.  Two things the reference cannot supply
on this image are provided as *user code*, exactly the way its API expects a user to provide them:

* the model/learner binding: the reference's PyTorch binding needs ``lightning`` + ``torchmetrics``, which are not
  installed and cannot be (no network), so ``TorchShimModel`` / ``TorchShimLearner`` subclass the reference's own
  ``P2PFLModel`` / ``Learner`` with the same ``state_dict <-> [ndarray]`` logic as its ``LightningModel``
  (p2pfl/learning/frameworks/pytorch/lightning_model.py:61-98) and a plain torch loop (Adam lr 1e-3, fresh per fit
  like a fresh Lightning Trainer, cross-entropy, accuracy);
* the CNN ``nn.Module`` (the reference ships only an MLP) - same architecture as the other arm, defined here.

The reference has no multi-GPU notion: all 8 nodes live in rank 0's process (its in-memory protocol cannot cross
processes) and their models are spread round-robin over the N visible GPUs; other ranks only join the barriers.
"""

from __future__ import annotations

import json
import os
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "baseline", "_ref")


def _unavailable(why: str) -> None:
    print(json.dumps({"impl": "reference", "unavailable": why}), flush=True)


def run_reference(args) -> None:
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    if not os.path.isdir(os.path.join(REF, "p2pfl")):
        if rank == 0:
            _unavailable("baseline/_ref/p2pfl is not installed (see DESIGN.md, 'Reference install')")
        return
    import torch

    multi = world > 1
    if multi:
        import torch.distributed as dist

        dist.init_process_group("gloo")
    try:
        if rank == 0:
            _run_rank0(args, world, torch)
    except Exception as e:  # noqa: BLE001
        if rank == 0:
            _unavailable(f"reference run failed: {type(e).__name__}: {e}"[:300])
    finally:
        if multi:
            dist.barrier()
            dist.destroy_process_group()


def _run_rank0(args, n_gpus: int, torch) -> None:
    import numpy as np
    from torch import nn

    sys.path.insert(0, REF)
    os.chdir(tempfile.mkdtemp(prefix="refarm-"))  # the reference writes logs/p2pfl.log into the CWD
    from p2pfl.communication.protocols.memory.memory_communication_protocol import InMemoryCommunicationProtocol
    from p2pfl.learning.aggregators.fedavg import FedAvg
    from p2pfl.learning.frameworks.learner import Learner
    from p2pfl.learning.frameworks.p2pfl_model import P2PFLModel
    from p2pfl.management.logger import logger
    from p2pfl.node import Node
    from p2pfl.settings import Settings

    sys.path.insert(0, ROOT)
    from bench import BATCH, DTYPE_NAMES, METRIC, SAMPLES_PER_NODE, TEST_PER_NODE, ClockSampler, common_config

    n_nodes = args.nodes
    has_cuda = torch.cuda.is_available()  # CPU fallback only for debugging the arm on a box without a GPU

    # ---- identical Settings values to the other arm (myfyp_b200.utils.utils.set_fast_settings)
    Settings.GRPC_TIMEOUT = 10
    Settings.HEARTBEAT_PERIOD = 1
    Settings.HEARTBEAT_TIMEOUT = 30
    Settings.GOSSIP_PERIOD = 0
    Settings.TTL = 10
    Settings.GOSSIP_MESSAGES_PER_PERIOD = 10_000_000
    Settings.AMOUNT_LAST_MESSAGES_SAVED = 10000
    Settings.GOSSIP_MODELS_PERIOD = 0.05
    Settings.GOSSIP_MODELS_PER_ROUND = 64
    Settings.GOSSIP_EXIT_ON_X_EQUAL_ROUNDS = 200
    Settings.TRAIN_SET_SIZE = n_nodes
    Settings.VOTE_TIMEOUT = 60
    Settings.AGGREGATION_TIMEOUT = 60
    Settings.WAIT_HEARTBEATS_CONVERGENCE = 0.0
    Settings.LOG_LEVEL = "WARNING"
    Settings.DISABLE_RAY = True
    logger.set_level("WARNING")

    class CNN(nn.Module):
        def __init__(self, in_channels=3, num_classes=10, width=32, image_size=32):
            super().__init__()
            w = width
            self.features = nn.Sequential(
                nn.Conv2d(in_channels, w, 3, padding=1), nn.ReLU(inplace=True), nn.MaxPool2d(2),
                nn.Conv2d(w, 2 * w, 3, padding=1), nn.ReLU(inplace=True), nn.MaxPool2d(2),
                nn.Conv2d(2 * w, 4 * w, 3, padding=1), nn.ReLU(inplace=True), nn.MaxPool2d(2),
            )
            s = image_size // 8
            self.classifier = nn.Sequential(nn.Flatten(), nn.Linear(4 * w * s * s, 256), nn.ReLU(inplace=True), nn.Linear(256, num_classes))

        def forward(self, x):
            return self.classifier(self.features(x.float()))

    class MLP(nn.Module):
        def __init__(self):
            super().__init__()
            self.layers = nn.ModuleList([nn.Linear(784, 256), nn.Linear(256, 128), nn.Linear(128, 10)])

        def forward(self, x):
            x = x.reshape(x.shape[0], -1).float()
            x = torch.relu(self.layers[0](x))
            x = torch.relu(self.layers[1](x))
            return self.layers[2](x)

    class TorchShimModel(P2PFLModel):
        def get_parameters(self):
            return [p.cpu().numpy() for _, p in self.model.state_dict().items()]

        def set_parameters(self, params):
            if isinstance(params, bytes):
                params, info = self.decode_parameters(params)
                self.additional_info.update(info)
            sd = self.model.state_dict()
            for (name, p), new in zip(sd.items(), params):
                if p.shape != new.shape:
                    raise ValueError("Not matching models")
                sd[name] = torch.tensor(new)
            self.model.load_state_dict(sd)

        def get_framework(self):
            return "pytorch"

    class ArrayData:
        def __init__(self, xtr, ytr, xte, yte):
            self.xtr, self.ytr, self.xte, self.yte = xtr, ytr, xte, yte

        def get_num_samples(self, train=True):
            return len(self.ytr if train else self.yte)

    class TorchShimLearner(Learner):
        data_on_host = False
        h2d = 0
        d2h = 0

        def __init__(self, model, data, self_addr="unknown-node", aggregator=None):
            super().__init__(model, data, self_addr, aggregator)
            self._stop = False
            self._cache = {}

        def _dev(self):
            return next(self.model.get_model().parameters()).device

        def _tensors(self, train):
            dev = self._dev()
            key = (train, TorchShimLearner.data_on_host)
            if key not in self._cache:
                x = torch.from_numpy(self.data.xtr if train else self.data.xte).float()
                y = torch.from_numpy(self.data.ytr if train else self.data.yte).long()
                self._cache[key] = (x.pin_memory() if has_cuda else x, y.pin_memory() if has_cuda else y) if TorchShimLearner.data_on_host else (x.to(dev), y.to(dev))
            x, y = self._cache[key]
            if TorchShimLearner.data_on_host:
                TorchShimLearner.h2d += x.numel() * 4 + y.numel() * 8
                x, y = x.to(dev, non_blocking=True), y.to(dev, non_blocking=True)
            return x, y

        def fit(self):
            self._stop = False
            module = self.model.get_model()
            dev = self._dev()
            x, y = self._tensors(True)
            opt = torch.optim.Adam(module.parameters(), lr=1e-3)
            module.train()
            losses = []
            for _ in range(self.epochs):
                perm = torch.randperm(x.shape[0], device=dev)
                for s in range(0, x.shape[0], BATCH):
                    if self._stop:
                        break
                    idx = perm[s : s + BATCH]
                    loss = torch.nn.functional.cross_entropy(module(x[idx]), y[idx])
                    opt.zero_grad()
                    loss.backward()
                    opt.step()
                    losses.append(loss.detach())
            if losses:
                vals = torch.stack(losses).tolist()
                TorchShimLearner.d2h += 4 * len(vals)
                for i, v in enumerate(vals):
                    logger.log_metric(self._self_addr, "train_loss", v, step=i)
            self.model.set_contribution([self._self_addr], self.data.get_num_samples(True))
            self.add_callback_info_to_model()
            return self.model

        def interrupt_fit(self):
            self._stop = True

        @torch.no_grad()
        def evaluate(self):
            module = self.model.get_model()
            x, y = self._tensors(False)
            module.eval()
            out = module(x)
            res = torch.stack([torch.nn.functional.cross_entropy(out, y), (out.argmax(1) == y).float().mean()]).tolist()
            TorchShimLearner.d2h += 8
            module.train()
            results = {"test_loss": res[0], "test_metric": res[1]}
            for k, v in results.items():
                logger.log_metric(self._self_addr, k, v)
            return results

        def get_framework(self):
            return "pytorch"

    # ---- same synthetic data recipe as the other arm (class-conditional Gaussian blobs, seed 0)
    shape = (28, 28) if args.model == "mlp" else (3, 32, 32)
    rng = np.random.default_rng(0)
    d = int(np.prod(shape))
    protos = rng.standard_normal((10, d)).astype(np.float32)

    def make(n):
        y = rng.integers(0, 10, size=n)
        x = rng.standard_normal((n, d)).astype(np.float32) * 0.8 + protos[y]
        return x.reshape((n, *shape)), y.astype(np.int64)

    xtr, ytr = make(SAMPLES_PER_NODE * n_nodes)
    xte, yte = make(TEST_PER_NODE * n_nodes)

    def sync():
        if has_cuda:
            for g in range(n_gpus):
                torch.cuda.synchronize(g)

    def measure(tag: str, host_mode: bool, warmup: int, steps: int, budget_s: float = 900.0):
        """One experiment of warmup+steps rounds on FRESH nodes; the last `steps` rounds are timed.

        (The reference cannot run two experiments on the same Node objects: its workflow object keeps
        `current_stage = RoundFinishedStage` after the first one, stages/workflows.py:37-50, so a second
        `set_start_learning` dies in `increase_round`.  Hence one experiment per measurement, and the warm-up
        rounds are the first rounds of that experiment.)"""
        TorchShimLearner.data_on_host = host_mode
        nodes = []
        for i in range(n_nodes):
            dev = torch.device("cuda", i * n_gpus // n_nodes) if has_cuda else torch.device("cpu")
            torch.manual_seed(i)
            module = (MLP() if args.model == "mlp" else CNN()).to(dev)
            tr = slice(i * SAMPLES_PER_NODE, (i + 1) * SAMPLES_PER_NODE)
            te = slice(i * TEST_PER_NODE, (i + 1) * TEST_PER_NODE)
            data = ArrayData(xtr[tr], ytr[tr], xte[te], yte[te])
            nodes.append(Node(TorchShimModel(module), data, address=f"{tag}-{i}", learner=TorchShimLearner, aggregator=FedAvg(), protocol=InMemoryCommunicationProtocol))
        for n in nodes:
            n.start()
        # Topology: a LINE (what the reference's own convergence test uses, test/node_test.py:92-96).  A fully connected
        # graph trips a race in the reference: two initialised neighbours deliver `init_model` at once, both pass the
        # `model_initialized_lock.locked()` check and the second `release()` raises (init_model_command.py:62-80); the
        # receiver then stops ITSELF (:92-97) and `InMemoryServer.stop()` wipes the process-wide registry
        # (memory_server.py:90-94), which takes the whole federation down.  Over a line nobody has two initialised
        # neighbours at the same time.  Partial-model gossip is unaffected: train-set members reach each other with
        # `create_connection=True` (train_stage.py:170-176), so the exchange is still all-to-all among the 8 nodes.
        for i in range(n_nodes - 1):
            nodes[i].connect(nodes[i + 1].addr)
            time.sleep(0.1)
        t_conv = time.time() + 30
        while any(len(n.get_neighbors(only_direct=False)) < n_nodes - 1 for n in nodes):
            if time.time() > t_conv:
                raise TimeoutError("reference heartbeat flood did not converge")
            time.sleep(0.1)
        sync()
        nodes[0].set_start_learning(rounds=warmup + steps, epochs=1)
        t_end = time.time() + budget_s

        def rounds_done(n):
            if n.learning_workflow.finished:
                return warmup + steps
            r = n.state.round
            return -1 if r is None else r

        while min(rounds_done(n) for n in nodes) < warmup:  # every node has finished the warm-up rounds
            if time.time() > t_end:
                raise TimeoutError("reference warm-up rounds did not finish")
            time.sleep(0.0005)
        TorchShimLearner.h2d = TorchShimLearner.d2h = 0
        if has_cuda:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
        t0 = time.perf_counter()
        while not all(n.learning_workflow.finished for n in nodes):
            if time.time() > t_end:
                raise TimeoutError(f"reference did not finish {warmup + steps} rounds in {budget_s:.0f} s")
            time.sleep(0.0005)
        if has_cuda:
            e1.record()
        sync()
        wall = time.perf_counter() - t0
        # many devices / threads: the host clock bounds the device span
        ms = max(e0.elapsed_time(e1), 1e3 * wall) if has_cuda else 1e3 * wall
        for n in nodes:
            try:
                n.stop()
            except Exception:  # noqa: BLE001
                pass
        time.sleep(0.5)
        return ms, wall, TorchShimLearner.h2d, TorchShimLearner.d2h

    sampler = ClockSampler()
    sampler.start()
    ms, _, _, _ = measure("ref", False, args.warmup, args.steps)
    clocks = sampler.stop(n_gpus)
    e2e = None
    if not args.no_e2e:
        _, w2, h2d, d2h = measure("refe2e", True, args.warmup, args.steps)
        e2e = {"value": args.steps / w2, "unit": "rounds/s", "h2d_bytes_per_step": int(h2d / args.steps), "d2h_bytes_per_step": int(d2h / args.steps), "ms_per_step": 1e3 * w2 / args.steps}
    value = args.steps / (ms / 1e3)
    out = {
        "metric": METRIC, "value": value, "unit": "rounds/s", "n_gpus": n_gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
        "higher_is_better": True, "scaling": "weak" if getattr(args, "nodes_per_gpu", 0) > 0 else "strong", "vs_baseline": None,
        # torch's defaults, read back from the process that trained: fp32 storage, cuDNN convolutions may use TF32 tensor cores
        # (torch.backends.cudnn.allow_tf32 = True), nn.Linear in plain fp32 (torch.backends.cuda.matmul.allow_tf32 = False)
        "dtype": DTYPE_NAMES["tf32"] if (args.model == "cnn" and torch.backends.cudnn.allow_tf32 and not torch.backends.cuda.matmul.allow_tf32) else "fp32 (torch default)",
        "data": "synthetic", "impl": "reference",
        "config": {
            **common_config(args, n_nodes), "topology": "line",
            "precision": f"torch defaults: fp32 storage, cudnn.allow_tf32={torch.backends.cudnn.allow_tf32}, cuda.matmul.allow_tf32={torch.backends.cuda.matmul.allow_tf32}",
            "topology_note": "a full mesh trips a reference race (benchmarks/reference_arm.py); partial-model gossip is all-to-all either way", "aggregator_impl": "reference NumPy FedAvg",
            "protocol": "in-memory", "learner": "plain-torch shim (lightning is not installable here)", "parallelism": f"{n_nodes} in-process nodes over {n_gpus} gpu(s)",
        },
        "clocks": clocks, "gpu_launches": 0,
    }
    if e2e is not None:
        out["e2e"] = e2e
    print(json.dumps(out), flush=True)
