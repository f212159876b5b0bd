"""The PyTorch learner: a hand-rolled, device-resident, CUDA-graph-captured training loop.

Replaces the reference's ``LightningLearner`` (pytorch/lightning_learner.py:49-152), which builds a new
``lightning.Trainer`` per ``fit`` and streams batch-size-1 batches through a DataLoader.  Here:

* the node's whole partition sits in HBM (``TensorExportStrategy``); a batch is an index gather;
* loss = fused log-softmax + NLL kernel (T3/T8); optimizer = ONE fused multi-tensor kernel over the flat slab (T6)
  that also applies FedProx / Scaffold terms and clears the gradients;
* a module exposing ``native_train_step`` (our tcgen05 models) bypasses autograd entirely;
* on CUDA one whole local epoch (gathers, forward, loss, backward, optimizer for every batch) is captured ONCE
  into a CUDA graph and replayed per epoch on the learner's own stream: with 4-8 nodes per GPU and 0.5-45 MB models
  the step is launch-bound, so a fit costs one graph launch instead of thousands of kernel launches, and the nodes
  of a GPU overlap on the device.  The Adam step counter lives on the device so replays keep exact bias corrections;
* per-step ``train_loss`` values stay on the device and reach the metric store with one D2H copy per fit;
  evaluation returns ``test_loss`` / ``test_metric`` (accuracy) like the reference
  (lightning_model.py:197-207).

Like the reference (a fresh Trainer => fresh optimizer per fit) the optimizer state is reset every ``fit`` unless
``persist_optimizer=True``.
"""

from __future__ import annotations

import threading
import zlib
from typing import Any, Dict, List, Optional

import torch

from myfyp_b200 import ops
from myfyp_b200.learning.dataset.p2pfl_dataset import TensorExportStrategy
from myfyp_b200.learning.learner import Learner
from myfyp_b200.learning.p2pfl_model import TorchP2PFLModel
from myfyp_b200.management.logger import logger
from myfyp_b200.management.trace import trace

_capture_lock = threading.Lock()  # one stream capture at a time per process


class TorchLearner(Learner):
    def __init__(
        self,
        model: TorchP2PFLModel,
        data,
        self_addr: str = "unknown-node",
        aggregator=None,
        batch_size: int = 32,
        lr: float = 1e-3,
        optimizer: str = "adam",
        momentum: float = 0.0,
        weight_decay: float = 0.0,
        persist_optimizer: bool = False,
        shuffle: bool = True,
        max_steps_per_epoch: Optional[int] = None,
        data_on_host: bool = False,
        use_cuda_graph: bool = True,
        extra_metrics: bool = False,
    ) -> None:
        super().__init__(model, data, self_addr, aggregator)
        self.batch_size, self.lr, self.optimizer = batch_size, lr, optimizer
        self.momentum, self.weight_decay = momentum, weight_decay
        self.persist_optimizer, self.shuffle = persist_optimizer, shuffle
        self.max_steps_per_epoch = max_steps_per_epoch
        self.use_cuda_graph = use_cuda_graph
        # extra_metrics: also report macro test_precision / test_recall / test_f1 (the fork author's MLP logs them)
        self.extra_metrics = extra_metrics
        # data_on_host: the partition stays in pinned host memory and is copied to the GPU at every fit /
        # evaluate (the end-to-end mode of bench.py); default is device-resident data
        self.data_on_host = data_on_host
        self.h2d_bytes = 0
        self.d2h_bytes = 0
        self.pre_fit_hook = None
        self._interrupt = False
        self._opt_state: Dict[str, Any] = {}
        self._data_cache: Dict[Any, Any] = {}
        self._static: Dict[Any, Any] = {}  # static device buffers the graphs read (x, y, perm, losses...)
        self._graphs: Dict[Any, Any] = {}
        self._stream: Optional[torch.cuda.Stream] = None
        self._global_step = 0
        self._rng = torch.Generator()
        self._rng.manual_seed(torch.initial_seed() ^ (zlib.crc32(self_addr.encode()) & 0xFFFFFFF))

    def get_framework(self) -> str:
        return "pytorch"

    # ------------------------------------------------------------------ data
    def _tensors(self, train: bool):
        dev = self.model.device
        host = self.data_on_host and dev.type == "cuda"
        key = (id(self.data), train, "host" if host else str(dev))
        if key not in self._data_cache:
            x, y = self.data.export(TensorExportStrategy, train=train, device="cpu" if host else dev)
            self._data_cache[key] = (x.pin_memory(), y.pin_memory()) if host else (x, y)
        x, y = self._data_cache[key]
        if host:
            # Every call consumes ONE host -> device copy of its inputs (pinned memory -> static device buffers).  The copies are
            # double-buffered and issued one call ahead on a copy stream: the copy that feeds call k+1 starts as soon as call k
            # has picked up its buffer, and overlaps call k's compute and the round's voting / aggregation (the copy engine is idle
            # otherwise) - what a prefetching data loader does.  Bytes per call are unchanged; the first call copies in line.
            self.h2d_bytes += x.numel() * x.element_size() + y.numel() * y.element_size()
            skey = ("xy", train)
            st = self._static.get(skey)
            cur = torch.cuda.current_stream(dev)
            if st is None:
                bufs = [(torch.empty_like(x, device=dev), torch.empty_like(y, device=dev)) for _ in range(2)]
                st = self._static[skey] = {"bufs": bufs, "next": 0, "event": None, "stream": torch.cuda.Stream(dev)}
            idx = st["next"]
            sx, sy = st["bufs"][idx]
            if st["event"] is None:  # nothing in flight for this buffer: copy now
                sx.copy_(x, non_blocking=True)
                sy.copy_(y, non_blocking=True)
            else:
                cur.wait_event(st["event"])
            # prefetch for the next call into the other buffer (its last reader, call k-1, has completed: fit / evaluate end with a
            # stream synchronisation)
            nx, ny = st["bufs"][idx ^ 1]
            with torch.cuda.stream(st["stream"]):
                nx.copy_(x, non_blocking=True)
                ny.copy_(y, non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(st["stream"])
            st["event"], st["next"] = ev, idx ^ 1
            return sx, sy
        return x, y

    def set_data(self, data) -> None:
        super().set_data(data)
        self._data_cache.clear()
        self._static.clear()
        self._graphs.clear()

    # ------------------------------------------------------------------ optimizer state
    def _state(self, n: int, device) -> Dict[str, Any]:
        st = self._opt_state
        if st.get("n") != n or st.get("device") != device:
            st.clear()
            st.update(n=n, device=device, m=torch.zeros(n, device=device), v=torch.zeros(n, device=device), step=0, dev_state=torch.zeros(4, device=device))
            self._graphs.clear()
        return st

    def _optimizer_step(self, st: Dict[str, Any], terms: Dict[str, Any], step_in_epoch: int, p16=None) -> None:
        model: TorchP2PFLModel = self.model  # type: ignore[assignment]
        n = model.layout.n_trainable
        if self.optimizer == "adam":
            # step = (steps of earlier epochs, on the device) + (index inside this epoch, baked): one launch per step
            ops.adam_step(
                model.slab[:n], model.grad_slab[:n], st["m"], st["v"], 0, lr=self.lr, weight_decay=self.weight_decay, dev_state=st["dev_state"], step_offset=step_in_epoch,
                p16=None if p16 is None else p16[:n], **terms,
            )
        else:
            ops.sgd_step(model.slab[:n], model.grad_slab[:n], st["m"], lr=self.lr, momentum=self.momentum, weight_decay=self.weight_decay, **terms)
            if p16 is not None:
                ops.cast_slab(model.slab[:n], p16[:n])

    # ------------------------------------------------------------------ one epoch (eager, or the body that gets captured)
    def _epoch_body(self, x, y, perm, losses, steps: int, st, terms) -> int:
        module = self.model.get_model()
        native = getattr(module, "native_train_step", None)
        nat = getattr(module, "native", None)
        p16 = getattr(nat, "p16", None)
        if nat is not None:
            nat.refresh_weights()  # bf16 mirror of the (possibly just aggregated) parameters; Adam keeps it current afterwards
        done = 0
        for s in range(steps):
            if self._interrupt:
                break
            idx = perm[s * self.batch_size : (s + 1) * self.batch_size]
            xb, yb = x.index_select(0, idx), y.index_select(0, idx)
            if native is not None:
                native(xb, yb, losses[s : s + 1])
            else:
                loss = ops.cross_entropy(module(xb), yb)
                loss.backward()
                losses[s : s + 1].copy_(loss.detach().reshape(1))
            self._optimizer_step(st, terms, s + 1, p16)
            done += 1
        if self.optimizer == "adam" and done:
            ops.adam_advance(st["dev_state"], done)
        return done

    def _graph_for(self, key, build):
        """Capture ``build()`` once into a CUDA graph (after an eager warm-up run of it on the learner's stream)."""
        if key in self._graphs:
            return self._graphs[key]
        model: TorchP2PFLModel = self.model  # type: ignore[assignment]
        stream = self._stream
        with _capture_lock:
            # warm-up: lazy inits (cuDNN plans, smem attributes, workspaces) must happen outside the capture;
            # it is a real step, so parameters / optimizer state are restored afterwards
            keep = (model.slab.clone(), model.grad_slab.clone(), self._opt_state["m"].clone(), self._opt_state["v"].clone(), self._opt_state["dev_state"].clone())
            build()
            stream.synchronize()
            model.slab.copy_(keep[0])
            model.grad_slab.copy_(keep[1])
            self._opt_state["m"].copy_(keep[2])
            self._opt_state["v"].copy_(keep[3])
            self._opt_state["dev_state"].copy_(keep[4])
            stream.synchronize()
            g = torch.cuda.CUDAGraph()
            before = ops.native.launches()
            with torch.cuda.graph(g, stream=stream, capture_error_mode="thread_local"):
                build()
            g.native_kernels = ops.native.launches() - before  # our kernels per replay (for launch accounting)
            ops.native.count_launch(-g.native_kernels)  # captured, not executed
        self._graphs[key] = g
        return g

    # ------------------------------------------------------------------ fit / evaluate
    def fit(self) -> TorchP2PFLModel:
        model: TorchP2PFLModel = self.model  # type: ignore[assignment]
        module = model.get_model()
        dev = model.device
        self._interrupt = False
        if self.pre_fit_hook is not None:
            self.pre_fit_hook()
        cuda = dev.type == "cuda"
        if cuda and self._stream is None:
            self._stream = torch.cuda.Stream(dev)
        ctx = torch.cuda.stream(self._stream) if cuda else _NullCtx()
        if cuda:
            self._stream.wait_stream(torch.cuda.current_stream(dev))  # parameters were written on the caller's stream
        trace(self._self_addr, "fit:enter")
        with ctx:
            x, y = self._tensors(train=True)
            n = x.shape[0]
            st = self._state(model.layout.n_trainable, dev)
            trace(self._self_addr, "fit:tensors")
            if not self.persist_optimizer:
                st["m"].zero_()
                st["v"].zero_()
                st["dev_state"].zero_()
                st["step"] = 0
            for cb in self.callbacks:
                cb.on_fit_start(self)
            terms: Dict[str, Any] = {}
            for cb in self.callbacks:
                terms.update(cb.optimizer_terms(self))
            steps_per_epoch = (n + self.batch_size - 1) // self.batch_size
            if self.max_steps_per_epoch is not None:
                steps_per_epoch = min(steps_per_epoch, self.max_steps_per_epoch)
            total = max(self.epochs * steps_per_epoch, 1)
            trace(self._self_addr, "fit:callbacks")
            all_losses = torch.zeros(total, device=dev)
            module.train()
            trace(self._self_addr, "fit:alloc")
            steps = 0
            graph_ok = cuda and self.use_cuda_graph
            grouped = self._cohort(module) if graph_ok and self.optimizer in ("adam", "sgd") else None
            if grouped is not None:
                steps, vals_grouped = self._fit_grouped(grouped, x, y, n, st, terms, steps_per_epoch)
            elif graph_ok:
                pkey = ("perm", n)
                if pkey not in self._static:
                    self._static[pkey] = (torch.arange(n, device=dev), torch.zeros(steps_per_epoch, device=dev))
                perm, losses = self._static[pkey]
                gkey = ("fit", n, steps_per_epoch, self.batch_size, self.optimizer, x.data_ptr(), tuple(sorted((k, v.data_ptr() if isinstance(v, torch.Tensor) else v) for k, v in terms.items())))
                for e in range(self.epochs):
                    if self._interrupt:
                        break
                    if self.shuffle:
                        # host-side generator: the device RNG must not be touched while ANOTHER node's thread is
                        # capturing its graph (PyTorch registers the CUDA generator with an active capture)
                        perm.copy_(torch.randperm(n, generator=self._rng))
                        trace(self._self_addr, "fit:perm")
                    g = self._graph_for(gkey, lambda: self._epoch_body(x, y, perm, losses, steps_per_epoch, st, terms))
                    trace(self._self_addr, "replay>")
                    g.replay()
                    trace(self._self_addr, "replay<")
                    all_losses[e * steps_per_epoch : (e + 1) * steps_per_epoch].copy_(losses)
                    steps += steps_per_epoch
                    ops.native.count_launch(g.native_kernels)
            else:
                for e in range(self.epochs):
                    perm = torch.randperm(n, generator=self._rng).to(dev) if self.shuffle else torch.arange(n, device=dev)
                    steps += self._epoch_body(x, y, perm, all_losses[e * steps_per_epoch :], steps_per_epoch, st, terms)
            for cb in self.callbacks:
                cb.on_fit_end(self, steps, self.lr)
            if grouped is not None:
                vals = vals_grouped
            else:
                vals = self._read_back(all_losses[:steps]) if steps else []  # one D2H copy per fit (synchronises the learner's stream)
        if cuda:
            self._stream.synchronize()
        if steps:
            self.d2h_bytes += 4 * steps
            logger.log_metric_series(self._self_addr, "train_loss", vals, self._global_step)
            self._global_step += steps
        if getattr(module, "native", None) is not None:
            model.sync_mirrors_from_slab()  # e.g. BatchNorm's integer batch counter, advanced on the device in its fp32 slab slot
        model.set_contribution([self._self_addr], self.data.get_num_samples(train=True))
        self.add_callback_info_to_model()
        return model

    def _read_back(self, t: torch.Tensor) -> List[float]:
        """Device -> host read of a small result vector WITHOUT holding the interpreter lock while the device finishes.

        ``Tensor.tolist()`` on a CUDA tensor waits for the stream inside the interpreter lock: with several nodes per process one
        node's wait for its own 20-50 ms fit froze every other node's thread (measured: the process made no progress for 60 %
        of a round).  An asynchronous copy into pinned memory plus ``Stream.synchronize()`` (which releases the lock) does not."""
        if not t.is_cuda:
            return t.tolist()
        n = t.numel()
        buf = self._static.get("host_readback")
        if buf is None or buf.numel() < n:
            buf = torch.empty(max(n, 256), dtype=torch.float32).pin_memory()
            self._static["host_readback"] = buf
        buf[:n].copy_(t.reshape(-1), non_blocking=True)
        torch.cuda.current_stream(t.device).synchronize()
        return buf[:n].tolist()

    # ------------------------------------------------------------------ node-batched TF32 path (models/grouped_cnn.py)
    def _cohort(self, module):
        from myfyp_b200.settings import Settings

        spec = getattr(module, "grouped_spec", None)
        if spec is None or getattr(Settings, "DISABLE_RAY", False):  # the reference's switch for its actor pool turns ours off too
            return None
        from myfyp_b200.models.grouped_cnn import CohortTrainer

        model: TorchP2PFLModel = self.model  # type: ignore[assignment]
        tr = CohortTrainer.get(spec, model.layout, model.device, self.batch_size)
        tr.register(self)
        return tr

    def _fit_grouped(self, trainer, x, y, n: int, st, terms, steps_per_epoch: int):
        """Hand this node's epoch(s) to the GPU's cohort trainer: it runs together with the other nodes that ask within the
        window, one graph replay per epoch for all of them.  Returns (steps done, per-step losses)."""
        model: TorchP2PFLModel = self.model  # type: ignore[assignment]
        dev = model.device
        pkey = ("gperm", n)
        if pkey not in self._static:
            self._static[pkey] = (torch.arange(n, device=dev), torch.zeros(max(steps_per_epoch, 2), device=dev), torch.empty(n, dtype=torch.int64).pin_memory())
        perm, losses, host_perm = self._static[pkey]

        def make_perm():
            torch.randperm(n, generator=self._rng, out=host_perm)
            return host_perm

        ev = torch.cuda.Event()
        ev.record(self._stream)
        args = dict(
            slab=model.slab, m=st["m"], v=st["v"], adam_state=st["dev_state"], x=x, y=y, perm=perm if self.shuffle else None, losses=losses, n_samples=n,
            lr=self.lr, weight_decay=self.weight_decay, optimizer=self.optimizer, momentum=self.momentum, epochs=self.epochs, max_steps=self.max_steps_per_epoch,
            shuffle=self.shuffle, make_perm=make_perm, ready=ev, **terms,
        )
        res = trainer.submit("fit", self, args)
        return res["steps"], res["losses"]

    def interrupt_fit(self) -> None:
        self._interrupt = True

    def _eval_body(self, x, y, acc, conf=None) -> None:
        module = self.model.get_model()
        native = getattr(module, "native_eval", None)
        if getattr(module, "native", None) is not None:
            module.native.refresh_weights()
        acc.zero_()
        if conf is not None:
            conf.zero_()
        bs = max(self.batch_size, 256)
        for s in range(0, x.shape[0], bs):
            xb, yb = x[s : s + bs], y[s : s + bs]
            if native is not None:
                native(xb, yb, acc)
                if conf is not None:
                    ops.confusion(module.native.logits(xb), yb, conf)
            else:
                logits = module(xb)
                ops.cross_entropy_eval(logits, yb, acc)
                if conf is not None:
                    ops.confusion(logits, yb, conf)

    @torch.no_grad()
    def evaluate(self) -> Dict[str, float]:
        model: TorchP2PFLModel = self.model  # type: ignore[assignment]
        module = model.get_model()
        dev = model.device
        cuda = dev.type == "cuda"
        if cuda and self._stream is None:
            self._stream = torch.cuda.Stream(dev)
        ctx = torch.cuda.stream(self._stream) if cuda else _NullCtx()
        if cuda:
            self._stream.wait_stream(torch.cuda.current_stream(dev))
        with ctx:
            try:
                x, y = self._tensors(train=False)
            except Exception:  # noqa: BLE001
                return {}
            if x.shape[0] == 0:
                return {}
            module.eval()
            if "acc" not in self._static:
                self._static["acc"] = torch.zeros(2, device=dev)  # [sum loss, correct]
            acc = self._static["acc"]
            conf = None
            if self.extra_metrics:
                ncls = int(y.max().item()) + 1 if "conf" not in self._static else self._static["conf"].shape[0]
                if "conf" not in self._static:
                    self._static["conf"] = torch.zeros(ncls, ncls, device=dev)
                conf = self._static["conf"]
            grouped = self._cohort(module) if cuda and self.use_cuda_graph and conf is None else None
            if grouped is not None:
                acc.zero_()
                ev = torch.cuda.Event()
                ev.record(self._stream)
                tot = grouped.submit("eval", self, {"slab": model.slab, "x": x, "y": y, "perm": None, "losses": acc, "n_samples": x.shape[0], "ready": ev})
            elif cuda and self.use_cuda_graph:
                self._state(model.layout.n_trainable, dev)
                g = self._graph_for(("eval", x.shape[0], x.data_ptr(), conf is not None), lambda: self._eval_body(x, y, acc, conf))
                g.replay()
                ops.native.count_launch(g.native_kernels)
            else:
                self._eval_body(x, y, acc, conf)
            module.train()
            if grouped is None:
                tot = self._read_back(acc)
        self.d2h_bytes += 8
        results = {"test_loss": tot[0] / x.shape[0], "test_metric": tot[1] / x.shape[0]}
        if conf is not None:
            results.update(ops.prf_from_confusion(conf))
        for k, v in results.items():
            logger.log_metric(self._self_addr, k, v)
        return results


class _NullCtx:
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False


# name used by scripts written against the reference
LightningLearner = TorchLearner
