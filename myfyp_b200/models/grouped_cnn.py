"""Node-batched TF32 training of the CNN: the federated nodes that share a GPU train in the SAME kernel launches.

The reference runs the learners of different simulated nodes in separate Ray actor processes
(p2pfl/learning/frameworks/simulation/actor_pool.py:101-128, virtual_learner.py:101-137).  A node's step here is ~4 GFLOP -
microseconds of tensor-core time behind a chain of latency-bound launches - so processes (or streams) per node leave the
GPU idle.  Instead every kernel of ``csrc/train_tf32.cu`` takes the node index from ``blockIdx.y``:

* :class:`CnnPlan` wraps one native plan = G *slots* (activation buffers, weight / gradient mirrors, tensor maps - all
  built once, in one workspace tensor) for one architecture and batch size;
* :class:`CohortTrainer` (one per GPU and architecture) is where learners meet: ``fit`` / ``evaluate`` calls that arrive
  within a short window are bound to slots and run as ONE CUDA-graph replay per epoch (13 launches per step for the whole
  cohort, whatever its size).  The first caller to arrive drives the batch, the others wait for their result - no
  extra thread, no process pool.

Precision: fp32 storage, TF32 tensor-core convolutions and fp32-class (3xTF32) linear layers - torch's defaults for the
same ``nn.Module`` (see the header of train_tf32.cu); the bf16 path of :mod:`native_cnn` stays available.
"""

from __future__ import annotations

import ctypes
import os
import threading
import time
import weakref
from typing import Any, Dict, List, Optional, Tuple

import torch
from torch import nn

from myfyp_b200.ops import native

MAX_CONV = 4
GROUP_SIZES = (1, 2, 4, 8, 16)

# ids of csrc/train_tf32.cu (tests drive single kernels and read slot buffers through these)
RUN_REFRESH, RUN_CONV0_FWD, RUN_FC1_FWD, RUN_HEAD_TRAIN, RUN_FC1_WGRAD, RUN_FC1_DGRAD, RUN_CONV0_WGRAD, RUN_ADAM, RUN_HEAD_EVAL = range(9)
RUN_CONV_FWD, RUN_CONV_WGRAD, RUN_CONV_DGRAD = 10, 20, 30
BUF_AHI, BUF_ALO, BUF_Z1, BUF_DHHI, BUF_DHLO, BUF_DPM0, BUF_WM, BUF_WLO, BUF_GM = range(1, 10)
BUF_ACT, BUF_IDX, BUF_DZ = 100, 110, 120


class _Desc(ctypes.Structure):
    _fields_ = [
        ("groups", ctypes.c_int), ("batch", ctypes.c_int), ("image", ctypes.c_int), ("in_ch", ctypes.c_int), ("n_conv", ctypes.c_int),
        ("conv_ch", ctypes.c_int * MAX_CONV), ("hidden", ctypes.c_int), ("classes", ctypes.c_int),
        ("off_conv_w", ctypes.c_int64 * MAX_CONV), ("off_conv_b", ctypes.c_int64 * MAX_CONV),
        ("off_fc1_w", ctypes.c_int64), ("off_fc1_b", ctypes.c_int64), ("off_fc2_w", ctypes.c_int64), ("off_fc2_b", ctypes.c_int64),
        ("slab_elems", ctypes.c_int64),
    ]


class _Bind(ctypes.Structure):
    _fields_ = [
        ("slab", ctypes.c_void_p), ("m", ctypes.c_void_p), ("v", ctypes.c_void_p), ("adam_state", ctypes.c_void_p),
        ("x", ctypes.c_void_p), ("y", ctypes.c_void_p), ("perm", ctypes.c_void_p), ("losses", ctypes.c_void_p),
        ("prox_ref", ctypes.c_void_p), ("cv_global", ctypes.c_void_p), ("cv_local", ctypes.c_void_p),
        ("prox_mu", ctypes.c_float), ("cv_scale", ctypes.c_float), ("lr", ctypes.c_float), ("weight_decay", ctypes.c_float),
        ("n_samples", ctypes.c_int), ("active", ctypes.c_int), ("sgd", ctypes.c_int), ("momentum", ctypes.c_float),
    ]


class CnnSpec:
    """Architecture of a ``models.cnn.CNN``-shaped module as the native plan needs it; ``None`` from :meth:`of` when the
    module has another shape (the caller then keeps the generic path)."""

    def __init__(self, in_ch: int, image: int, conv_ch: Tuple[int, ...], hidden: int, classes: int, names: Dict[str, str]) -> None:
        self.in_ch, self.image, self.conv_ch, self.hidden, self.classes, self.names = in_ch, image, tuple(conv_ch), hidden, classes, names

    @staticmethod
    def of(module: nn.Module, image: int) -> Optional["CnnSpec"]:
        feats = getattr(module, "features", None)
        cls = getattr(module, "classifier", None)
        if feats is None or cls is None:
            return None
        convs, names, i, fl = [], {}, 0, list(feats)
        while i < len(fl):
            m = fl[i]
            if not isinstance(m, nn.Conv2d):
                return None
            ok = m.kernel_size == (3, 3) and m.stride == (1, 1) and m.padding == (1, 1) and m.bias is not None and m.groups == 1
            if not ok or i + 2 >= len(fl) or not isinstance(fl[i + 1], nn.ReLU) or not isinstance(fl[i + 2], nn.MaxPool2d) or fl[i + 2].kernel_size not in (2, (2, 2)):
                return None
            names[f"conv{len(convs)}"] = f"features.{i}"
            convs.append(m)
            i += 3
        lin = [(j, m) for j, m in enumerate(cls) if isinstance(m, nn.Linear)]
        if len(convs) < 2 or len(convs) > MAX_CONV or len(lin) != 2 or any(not isinstance(m, (nn.Flatten, nn.Linear, nn.ReLU)) for m in cls):
            return None
        for a, b in zip(convs[:-1], convs[1:]):
            if a.out_channels != b.in_channels:
                return None
        names["fc1"], names["fc2"] = f"classifier.{lin[0][0]}", f"classifier.{lin[1][0]}"
        fh = image >> len(convs)
        if lin[0][1].in_features != fh * fh * convs[-1].out_channels or lin[0][1].out_features != lin[1][1].in_features:
            return None
        return CnnSpec(convs[0].in_channels, image, tuple(c.out_channels for c in convs), lin[0][1].out_features, lin[1][1].out_features, names)

    def desc(self, layout, groups: int, batch: int) -> _Desc:
        d = _Desc()
        d.groups, d.batch, d.image, d.in_ch, d.n_conv = groups, batch, self.image, self.in_ch, len(self.conv_ch)
        for i, c in enumerate(self.conv_ch):
            d.conv_ch[i] = c
            d.off_conv_w[i] = layout.entries[self.names[f"conv{i}"] + ".weight"].offset
            d.off_conv_b[i] = layout.entries[self.names[f"conv{i}"] + ".bias"].offset
        d.hidden, d.classes = self.hidden, self.classes
        d.off_fc1_w = layout.entries[self.names["fc1"] + ".weight"].offset
        d.off_fc1_b = layout.entries[self.names["fc1"] + ".bias"].offset
        d.off_fc2_w = layout.entries[self.names["fc2"] + ".weight"].offset
        d.off_fc2_b = layout.entries[self.names["fc2"] + ".bias"].offset
        d.slab_elems = layout.n_trainable
        return d

    def key(self, layout) -> tuple:
        return (self.in_ch, self.image, self.conv_ch, self.hidden, self.classes, layout.hash)


def _lib():
    lib = native.load()
    if not getattr(lib, "_cnn_sigs", False):
        lib.b2_cnn_workspace_bytes.restype = ctypes.c_int64
        lib.b2_cnn_plan_create.restype = ctypes.c_void_p
        lib.b2_cnn_plan_create.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64, ctypes.c_void_p]
        for f in (lib.b2_cnn_plan_destroy, ):
            f.argtypes = [ctypes.c_void_p]
        lib.b2_cnn_plan_bind.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
        lib.b2_cnn_plan_refresh.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        for f in (lib.b2_cnn_plan_train_step, lib.b2_cnn_plan_eval_step, lib.b2_cnn_plan_advance):
            f.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p]
        lib.b2_cnn_plan_run.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p]
        lib.b2_cnn_plan_query.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p]
        lib._cnn_sigs = True
    return lib


class CnnPlan:
    """G slots of the native TF32 training step for one architecture / batch size on one GPU."""

    def __init__(self, spec: CnnSpec, layout, device: torch.device, groups: int, batch: int) -> None:
        self.spec, self.device, self.groups, self.batch = spec, device, groups, batch
        lib = _lib()
        self._desc = spec.desc(layout, groups, batch)
        nbytes = lib.b2_cnn_workspace_bytes(ctypes.byref(self._desc))
        if nbytes <= 0:
            native.check(1, "b2_cnn_workspace_bytes")
        with native.device_guard(device):
            self.workspace = torch.empty(nbytes + 1024, dtype=torch.uint8, device=device)
            base = (self.workspace.data_ptr() + 1023) // 1024 * 1024
            self._h = lib.b2_cnn_plan_create(ctypes.byref(self._desc), ctypes.c_void_p(base), ctypes.c_int64(nbytes), native.stream_ptr(device))
        if not self._h:
            native.check(1, "b2_cnn_plan_create")
        self._binds = (_Bind * groups)()

    def close(self) -> None:
        if getattr(self, "_h", None):
            _lib().b2_cnn_plan_destroy(ctypes.c_void_p(self._h))
            self._h = None

    def __del__(self) -> None:  # pragma: no cover - interpreter shutdown order
        try:
            self.close()
        except Exception:  # noqa: BLE001
            pass

    # ------------------------------------------------------------------ slots
    def bind(self, slots: List[Optional[dict]]) -> None:
        """``slots[i]``: tensors / scalars of the node in slot i (``None`` = slot idle); copied to the device bind table on the
        current stream."""
        p = native.ptr
        for i in range(self.groups):
            b, s = self._binds[i], (slots[i] if i < len(slots) else None)
            if s is None:
                b.active = 0
                continue
            b.slab, b.m, b.v, b.adam_state = p(s["slab"]), p(s.get("m")), p(s.get("v")), p(s.get("adam_state"))
            b.x, b.y, b.perm, b.losses = p(s["x"]), p(s["y"]), p(s.get("perm")), p(s["losses"])
            b.prox_ref, b.cv_global, b.cv_local = p(s.get("prox_ref")), p(s.get("cv_global")), p(s.get("cv_local"))
            b.prox_mu, b.cv_scale = float(s.get("prox_mu", 0.0)) if s.get("prox_ref") is not None else 0.0, float(s.get("cv_scale", 1.0))
            b.lr, b.weight_decay = float(s.get("lr", 1e-3)), float(s.get("weight_decay", 0.0))
            b.n_samples, b.active = int(s["n_samples"]), 1
            b.sgd, b.momentum = (1 if s.get("optimizer", "adam") == "sgd" else 0), float(s.get("momentum", 0.0))
        with native.device_guard(self.device):
            native.check(_lib().b2_cnn_plan_bind(self._h, ctypes.byref(self._binds), native.stream_ptr(self.device)), "b2_cnn_plan_bind")

    def _call(self, fn, name: str, *args) -> int:
        with native.device_guard(self.device):
            rc = fn(self._h, *args, native.stream_ptr(self.device))
        return rc

    def refresh(self) -> None:
        native.check(self._call(_lib().b2_cnn_plan_refresh, "refresh"), "b2_cnn_plan_refresh")
        native.count_launch()

    def train_step(self, step: int) -> int:
        n = self._call(_lib().b2_cnn_plan_train_step, "train", ctypes.c_int(step))
        if n < 0:
            native.check(1, "b2_cnn_plan_train_step")
        native.count_launch(n)
        return n

    def eval_step(self, step: int) -> int:
        n = self._call(_lib().b2_cnn_plan_eval_step, "eval", ctypes.c_int(step))
        if n < 0:
            native.check(1, "b2_cnn_plan_eval_step")
        native.count_launch(n)
        return n

    def advance(self, steps: int) -> None:
        native.check(self._call(_lib().b2_cnn_plan_advance, "advance", ctypes.c_int(steps)), "b2_cnn_plan_advance")
        native.count_launch()

    def run(self, which: int, step: int = 0) -> None:
        native.check(self._call(_lib().b2_cnn_plan_run, "run", ctypes.c_int(which), ctypes.c_int(step)), f"b2_cnn_plan_run({which})")
        native.count_launch()

    def buffer(self, slot: int, buf: int, dtype=torch.float32) -> torch.Tensor:
        """View of a slot buffer inside the workspace (tests / diagnostics)."""
        ptr, nbytes = ctypes.c_void_p(), ctypes.c_int64()
        native.check(_lib().b2_cnn_plan_query(self._h, slot, buf, ctypes.byref(ptr), ctypes.byref(nbytes)), "b2_cnn_plan_query")
        off = ptr.value - self.workspace.data_ptr()
        return self.workspace[off : off + nbytes.value].view(dtype)


# ====================================================================================================== cohort trainer
def _logger():
    from myfyp_b200.management.logger import logger

    return logger


class _Job:
    __slots__ = ("kind", "learner", "args", "done", "result", "error", "t_in")

    def __init__(self, kind: str, learner, args: dict) -> None:
        self.kind, self.learner, self.args = kind, learner, args
        self.done, self.result, self.error = False, None, None
        self.t_in = time.perf_counter()


class CohortTrainer:
    """Meeting point of the learners of one GPU that share an architecture.  ``fit`` / ``evaluate`` requests arriving within
    ``window`` seconds (or as soon as every registered learner has asked) run together."""

    _instances: Dict[tuple, "CohortTrainer"] = {}
    _lock = threading.Lock()

    @classmethod
    def get(cls, spec: CnnSpec, layout, device: torch.device, batch: int) -> "CohortTrainer":
        key = (str(device), batch) + spec.key(layout)
        with cls._lock:
            tr = cls._instances.get(key)
            if tr is None:
                tr = cls._instances[key] = CohortTrainer(spec, layout, device, batch)
            return tr

    @classmethod
    def reset(cls) -> None:
        with cls._lock:
            for tr in cls._instances.values():
                for p in tr._plans.values():
                    p.close()
            cls._instances.clear()

    def __init__(self, spec: CnnSpec, layout, device: torch.device, batch: int) -> None:
        from myfyp_b200.settings import Settings

        self.spec, self.layout, self.device, self.batch = spec, layout, device, batch
        self.window = float(getattr(Settings, "COHORT_WINDOW", 0.004))
        self.max_group = int(getattr(Settings, "COHORT_MAX", 8))
        self._plans: Dict[int, CnnPlan] = {}
        self._graphs: Dict[tuple, Any] = {}
        self._cv = threading.Condition()
        self._pending: List[_Job] = []
        self._leader = False
        self._members: "weakref.WeakSet" = weakref.WeakSet()
        self._stream: Optional[torch.cuda.Stream] = None
        self.batches = 0     # statistics: cohort launches and their total size
        self.batched = 0

    def register(self, learner) -> None:
        self._members.add(learner)

    # ------------------------------------------------------------------ plans / graphs
    def plan(self, n: int) -> CnnPlan:
        g = next(s for s in GROUP_SIZES if s >= n)
        if g not in self._plans:
            self._plans[g] = CnnPlan(self.spec, self.layout, self.device, g, self.batch)
        return self._plans[g]

    def _graph(self, plan: CnnPlan, kind: str, steps: int):
        key = (plan.groups, kind, steps)
        g = self._graphs.get(key)
        if g is None:
            # warm-up with every slot idle: all kernels return at once, but lazy one-time set-up (shared-memory attributes,
            # module load) happens outside the capture
            plan.bind([None] * plan.groups)
            plan.refresh()
            (plan.train_step if kind == "fit" else plan.eval_step)(0)
            self._stream.synchronize()
            g = torch.cuda.CUDAGraph()
            before = native.launches()
            with torch.cuda.graph(g, stream=self._stream, capture_error_mode="thread_local"):
                for s in range(steps):
                    (plan.train_step if kind == "fit" else plan.eval_step)(s)
                if kind == "fit":
                    plan.advance(steps)
            g.native_kernels = native.launches() - before
            native.count_launch(-g.native_kernels)
            self._graphs[key] = g
        return g

    # ------------------------------------------------------------------ submit
    def submit(self, kind: str, learner, args: dict):
        job = _Job(kind, learner, args)
        with self._cv:
            self._pending.append(job)
            self._cv.notify_all()
            while not job.done:
                if self._leader:
                    self._cv.wait()
                    continue
                self._leader = True
                batch: List[_Job] = []
                try:
                    deadline = time.monotonic() + self.window
                    while True:
                        same = [j for j in self._pending if j.kind == self._pending[0].kind]
                        want = min(self.max_group, max(1, len(self._members)))
                        left = deadline - time.monotonic()
                        if len(same) >= want or left <= 0:
                            break
                        self._cv.wait(left)
                    kind0 = self._pending[0].kind
                    batch = [j for j in self._pending if j.kind == kind0][: self.max_group]
                    for j in batch:
                        self._pending.remove(j)
                    self._cv.release()
                    try:
                        self._execute(kind0, batch)
                    except BaseException as e:  # noqa: BLE001 - every waiter must be released
                        for j in batch:
                            j.error = e
                    finally:
                        self._cv.acquire()
                finally:
                    for j in batch:
                        j.done = True
                    self._leader = False
                    self._cv.notify_all()
        if job.error is not None:
            raise job.error
        return job.result

    def _pinned(self, rows: int, cols: int) -> torch.Tensor:
        """Pinned read-back buffer: an asynchronous copy + ``Stream.synchronize()`` waits WITHOUT the interpreter lock, a plain
        ``.cpu()`` / ``.tolist()`` would freeze every other node thread of the process for the length of the epoch."""
        buf = getattr(self, "_host", None)
        if buf is None or buf.shape[0] < rows or buf.shape[1] < cols:
            buf = torch.empty(max(rows, self.max_group), max(cols, 256), dtype=torch.float32).pin_memory()
            self._host = buf
        return buf

    def _profile_kernels(self, plan: CnnPlan, st) -> None:
        """``B2_COHORT_PROFILE=1``: CUDA-event time of every kernel of the step on the slots bound right now (the live data of a
        federation, not the synthetic tensors of tools/t32_step_profile.py).  10 launches each; logged as ``cohortk/<name>`` (us)."""
        seq = [("conv0_fwd", RUN_CONV0_FWD), ("conv1_fwd", RUN_CONV_FWD + 1), ("conv2_fwd", RUN_CONV_FWD + 2), ("fc1_fwd", RUN_FC1_FWD), ("head", RUN_HEAD_TRAIN),
               ("fc1_wgrad", RUN_FC1_WGRAD), ("fc1_dgrad", RUN_FC1_DGRAD), ("conv2_wgrad", RUN_CONV_WGRAD + 2), ("conv2_dgrad", RUN_CONV_DGRAD + 2),
               ("conv1_wgrad", RUN_CONV_WGRAD + 1), ("conv1_dgrad", RUN_CONV_DGRAD + 1), ("conv0_wgrad", RUN_CONV0_WGRAD), ("adam", RUN_ADAM)]
        for name, which in seq:
            plan.run(which, 1)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(st)
            for _ in range(10):
                plan.run(which, 1)
            e1.record(st)
            st.synchronize()
            _logger().log_device_time(f"cohortk/{name}", 100.0 * e0.elapsed_time(e1))
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st)
        for s in range(10):
            plan.train_step(s)
        e1.record(st)
        st.synchronize()
        _logger().log_device_time("cohortk/step_eager", 100.0 * e0.elapsed_time(e1))

    # ------------------------------------------------------------------ one cohort
    def _execute(self, kind: str, batch: List[_Job]) -> None:
        dev = self.device
        if self._stream is None:
            self._stream = torch.cuda.Stream(dev)
        st = self._stream
        plan = self.plan(len(batch))
        self.batches += 1
        self.batched += len(batch)
        t0 = time.perf_counter()
        marks = [t0]

        def lap(name: str) -> None:  # wall-time split of a cohort, next to the stage laps (logger.get_device_times())
            marks.append(time.perf_counter())
            _logger().log_device_time(f"cohort/{kind}_{name}", 1e3 * (marks[-1] - marks[-2]))

        _logger().log_device_time(f"cohort/{kind}_gather", 1e3 * (t0 - min(j.t_in for j in batch)))  # first request -> everybody here
        with torch.cuda.device(dev), torch.cuda.stream(st):
            for j in batch:  # inputs were produced on the learners' own streams
                ev = j.args.get("ready")
                if ev is not None:
                    st.wait_event(ev)
            B = self.batch
            steps = max((int(j.args["n_samples"]) + B - 1) // B for j in batch)
            if kind == "fit":
                steps = min([steps] + [int(j.args["max_steps"]) for j in batch if j.args.get("max_steps")])
                epochs = max(int(j.args.get("epochs", 1)) for j in batch)
                slots = []
                for j in batch:
                    a = j.args
                    a["losses"].zero_()
                    slots.append(a)
                graph = self._graph(plan, "fit", steps)
                plan.bind(slots + [None] * (plan.groups - len(slots)))
                plan.refresh()
                host = self._pinned(len(batch), epochs * steps)
                for e in range(epochs):
                    live = [j for j in batch if e < int(j.args.get("epochs", 1))]
                    if len(live) != len(batch):
                        plan.bind([j.args if j in live else None for j in batch] + [None] * (plan.groups - len(batch)))
                    if e:
                        st.synchronize()  # the pinned permutation buffers are rewritten below
                    for j in live:
                        a = j.args
                        if a.get("perm") is not None and a.get("shuffle", True):
                            a["perm"].copy_(a["make_perm"](), non_blocking=True)
                        if e:
                            a["losses"].zero_()
                    if e == 0 and os.environ.get("B2_COHORT_PROFILE"):
                        keep = [(a[k], a[k].clone()) for a in slots for k in ("slab", "m", "v", "adam_state")]
                        self._profile_kernels(plan, st)  # debugging aid; the training state is put back afterwards
                        for t, c in keep:
                            t.copy_(c)
                        plan.refresh()
                    if e == 0:
                        lap("setup")
                        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                        ev0.record(st)
                    graph.replay()
                    if e == 0:
                        ev1.record(st)
                        lap("launch")
                    native.count_launch(graph.native_kernels)
                    for k, j in enumerate(batch):
                        if j in live:
                            host[k, e * steps : (e + 1) * steps].copy_(j.args["losses"][:steps], non_blocking=True)
                st.synchronize()  # releases the interpreter lock while the device works
                lap("device")
                _logger().log_device_time("cohort/fit_graph_on_device", ev0.elapsed_time(ev1))  # CUDA-event time of the first epoch's graph
                for k, j in enumerate(batch):
                    n_mine = min(steps, (int(j.args["n_samples"]) + B - 1) // B)
                    ep = int(j.args.get("epochs", 1))
                    vals = host[k, : ep * steps].view(ep, steps)[:, :n_mine].reshape(-1).tolist()
                    j.result = {"losses": vals, "steps": len(vals)}
                lap("post")
            else:
                for j in batch:
                    j.args["losses"].zero_()
                graph = self._graph(plan, "eval", steps)
                plan.bind([j.args for j in batch] + [None] * (plan.groups - len(batch)))
                plan.refresh()
                lap("setup")
                graph.replay()
                native.count_launch(graph.native_kernels)
                host = self._pinned(len(batch), 2)
                for k, j in enumerate(batch):
                    host[k, :2].copy_(j.args["losses"][:2], non_blocking=True)
                st.synchronize()
                lap("device")
                for k, j in enumerate(batch):
                    j.result = host[k, :2].tolist()


def attach_grouped_cnn(model, image_size: int = 32, batch_size: Optional[int] = None) -> Optional[CnnSpec]:
    """Mark a CUDA ``TorchP2PFLModel`` around a CNN for node-batched TF32 training; returns ``None`` (and changes nothing)
    when the architecture is outside what the native plan covers."""
    if model.device.type != "cuda":
        raise RuntimeError("the grouped TF32 path needs a CUDA device")
    module = model.get_model()
    spec = CnnSpec.of(module, image_size)
    if spec is None or image_size != 32 or spec.in_ch not in (1, 3) or any(c % 32 for c in spec.conv_ch) or spec.conv_ch[0] > 64 or spec.hidden % 64 or not (2 <= spec.classes <= 16):
        return None
    fh = image_size >> len(spec.conv_ch)
    if fh < 1 or (fh * fh * spec.conv_ch[-1]) % 64 or image_size >> (len(spec.conv_ch) - 1) < 4:
        return None
    module.grouped_spec = spec
    return spec
