"""The PyTorch learner: a hand-rolled, device-resident training loop.

Replaces the reference's ``LightningLearner`` (pytorch/lightning_learner.py:49-152), which builds a
new ``lightning.Trainer`` per ``fit`` and streams batch-size-1 batches through a DataLoader.  Here:

* the node's whole partition sits in HBM (``TensorExportStrategy``); a batch is an index slice;
* loss = fused log-softmax + NLL kernel (T3/T8); optimizer = ONE fused multi-tensor kernel over the
  flat slab (T6) that also applies FedProx / Scaffold terms and clears the gradients;
* a module exposing ``native_train_step`` (our tcgen05 models) bypasses autograd entirely;
* per-step ``train_loss`` values stay on the device and reach the metric store with one D2H copy
  per fit; evaluation returns ``test_loss`` / ``test_metric`` (accuracy) like the reference
  (lightning_model.py:197-207).

Like the reference (a fresh Trainer => fresh optimizer per fit) the optimizer state is reset every
``fit`` unless ``persist_optimizer=True``.
"""

from __future__ import annotations

from typing import Any, Dict, Optional

import torch

from myfyp_b200 import ops
from myfyp_b200.learning.dataset.p2pfl_dataset import TensorExportStrategy
from myfyp_b200.learning.learner import Learner
from myfyp_b200.learning.p2pfl_model import TorchP2PFLModel
from myfyp_b200.management.logger import logger


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
    ) -> None:
        super().__init__(model, data, self_addr, aggregator)
        self.batch_size, self.lr, self.optimizer = batch_size, lr, optimizer
        self.momentum, self.weight_decay = momentum, weight_decay
        self.persist_optimizer, self.shuffle = persist_optimizer, shuffle
        self.max_steps_per_epoch = max_steps_per_epoch
        # data_on_host: the partition stays in pinned host memory and is copied to the GPU at every fit /
        # evaluate (the end-to-end mode of bench.py); default is device-resident data
        self.data_on_host = data_on_host
        self.h2d_bytes = 0
        self.d2h_bytes = 0
        self.pre_fit_hook = None
        self._interrupt = False
        self._opt_state: Dict[str, Any] = {}
        self._data_cache: Dict[Any, Any] = {}
        self._global_step = 0

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
        if host:  # this call's inputs travel host -> device now
            self.h2d_bytes += x.numel() * x.element_size() + y.numel() * y.element_size()
            x, y = x.to(dev, non_blocking=True), y.to(dev, non_blocking=True)
        return x, y

    def set_data(self, data) -> None:
        super().set_data(data)
        self._data_cache.clear()

    # ------------------------------------------------------------------ optimizer state
    def _state(self, n: int, device) -> Dict[str, Any]:
        st = self._opt_state
        if st.get("n") != n or st.get("device") != device:
            st.clear()
            st.update(n=n, device=device, m=torch.zeros(n, device=device), v=torch.zeros(n, device=device), step=0)
        return st

    def _optimizer_step(self, st: Dict[str, Any], terms: Dict[str, Any]) -> None:
        model: TorchP2PFLModel = self.model  # type: ignore[assignment]
        n = model.layout.n_trainable
        st["step"] += 1
        if self.optimizer == "adam":
            ops.adam_step(model.slab[:n], model.grad_slab[:n], st["m"], st["v"], st["step"], lr=self.lr, weight_decay=self.weight_decay, **terms)
        else:
            ops.sgd_step(model.slab[:n], model.grad_slab[:n], st["m"], lr=self.lr, momentum=self.momentum, weight_decay=self.weight_decay, **terms)

    # ------------------------------------------------------------------ fit / evaluate
    def fit(self) -> TorchP2PFLModel:
        model: TorchP2PFLModel = self.model  # type: ignore[assignment]
        module = model.get_model()
        self._interrupt = False
        if self.pre_fit_hook is not None:
            self.pre_fit_hook()
        x, y = self._tensors(train=True)
        n = x.shape[0]
        st = self._state(model.layout.n_trainable, model.device)
        if not self.persist_optimizer:
            st["m"].zero_()
            st["v"].zero_()
            st["step"] = 0
        for cb in self.callbacks:
            cb.on_fit_start(self)
        terms: Dict[str, Any] = {}
        for cb in self.callbacks:
            terms.update(cb.optimizer_terms(self))
        steps_per_epoch = (n + self.batch_size - 1) // self.batch_size
        if self.max_steps_per_epoch is not None:
            steps_per_epoch = min(steps_per_epoch, self.max_steps_per_epoch)
        losses = torch.zeros(max(self.epochs * steps_per_epoch, 1), device=model.device)
        native = getattr(module, "native_train_step", None)
        module.train()
        steps = 0
        for _ in range(self.epochs):
            perm = torch.randperm(n, device=model.device) if self.shuffle else torch.arange(n, device=model.device)
            for s in range(steps_per_epoch):
                if self._interrupt:
                    break
                idx = perm[s * self.batch_size : (s + 1) * self.batch_size]
                xb, yb = x[idx], y[idx]
                if native is not None:
                    native(xb, yb, losses[steps : steps + 1])
                else:
                    loss = ops.cross_entropy(module(xb), yb)
                    loss.backward()
                    losses[steps] = loss.detach()
                self._optimizer_step(st, terms)
                steps += 1
        for cb in self.callbacks:
            cb.on_fit_end(self, steps, self.lr)
        if steps:
            self.d2h_bytes += 4 * steps
            for i, v in enumerate(losses[:steps].tolist()):  # one D2H copy per fit
                logger.log_metric(self._self_addr, "train_loss", v, step=self._global_step + i)
            self._global_step += steps
        model.set_contribution([self._self_addr], self.data.get_num_samples(train=True))
        self.add_callback_info_to_model()
        return model

    def interrupt_fit(self) -> None:
        self._interrupt = True

    @torch.no_grad()
    def evaluate(self) -> Dict[str, float]:
        model: TorchP2PFLModel = self.model  # type: ignore[assignment]
        module = model.get_model()
        try:
            x, y = self._tensors(train=False)
        except Exception:  # noqa: BLE001
            return {}
        if x.shape[0] == 0:
            return {}
        module.eval()
        native = getattr(module, "native_eval", None)
        acc = torch.zeros(2, device=model.device)  # [sum loss, correct]
        bs = max(self.batch_size, 256)
        for s in range(0, x.shape[0], bs):
            xb, yb = x[s : s + bs], y[s : s + bs]
            if native is not None:
                native(xb, yb, acc)
            else:
                ops.cross_entropy_eval(module(xb), yb, acc)
        module.train()
        tot = acc.tolist()
        self.d2h_bytes += 8
        results = {"test_loss": tot[0] / x.shape[0], "test_metric": tot[1] / x.shape[0]}
        for k, v in results.items():
            logger.log_metric(self._self_addr, k, v)
        return results


# name used by scripts written against the reference
LightningLearner = TorchLearner
