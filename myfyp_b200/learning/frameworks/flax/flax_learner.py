"""``FlaxLearner`` (reference: flax/flax_learner.py:40-181): optax Adam, a jitted batched step (the reference differentiates
sample by sample without jit)."""

from __future__ import annotations

from typing import Dict

import jax
import jax.numpy as jnp
import optax

from myfyp_b200.learning.frameworks.flax.flax_dataset import FlaxExportStrategy
from myfyp_b200.learning.learner import Learner
from myfyp_b200.management.logger import logger


class FlaxLearner(Learner):
    def __init__(self, model, data, self_addr: str = "unknown-node", aggregator=None, batch_size: int = 32, lr: float = 1e-3, **kwargs) -> None:
        super().__init__(model, data, self_addr, aggregator)
        self.batch_size = batch_size
        self.tx = optax.adam(lr)
        self.opt_state = self.tx.init(self.model.model_params)
        self._stop = False
        self._step = 0
        module = self.model.get_model()

        def loss_fn(params, x, y):
            logits = module.apply({"params": params}, x)
            return optax.softmax_cross_entropy_with_integer_labels(logits, y).mean(), logits

        @jax.jit
        def train_step(params, opt_state, x, y):
            (loss, _), grads = jax.value_and_grad(loss_fn, has_aux=True)(params, x, y)
            updates, opt_state = self.tx.update(grads, opt_state, params)
            return optax.apply_updates(params, updates), opt_state, loss

        @jax.jit
        def eval_step(params, x, y):
            loss, logits = loss_fn(params, x, y)
            return loss * x.shape[0], jnp.sum(jnp.argmax(logits, -1) == y)

        self._train_step, self._eval_step = train_step, eval_step

    def get_framework(self) -> str:
        return "flax"

    def fit(self):
        self._stop = False
        batches = self.data.export(FlaxExportStrategy, train=True, batch_size=self.batch_size)
        params = self.model.model_params
        for _ in range(self.epochs):
            for x, y in batches:
                if self._stop:
                    break
                params, self.opt_state, loss = self._train_step(params, self.opt_state, jnp.asarray(x), jnp.asarray(y))
                logger.log_metric(self._self_addr, "train_loss", float(loss), step=self._step)
                self._step += 1
        self.model.model_params = params
        self.model.set_contribution([self._self_addr], self.data.get_num_samples(train=True))
        self.add_callback_info_to_model()
        return self.model

    def interrupt_fit(self) -> None:
        self._stop = True

    def evaluate(self) -> Dict[str, float]:
        batches = self.data.export(FlaxExportStrategy, train=False, batch_size=max(self.batch_size, 256))
        tot_loss, correct, n = 0.0, 0, 0
        for x, y in batches:
            l, c = self._eval_step(self.model.model_params, jnp.asarray(x), jnp.asarray(y))
            tot_loss, correct, n = tot_loss + float(l), correct + int(c), n + len(y)
        if n == 0:
            return {}
        out = {"test_loss": tot_loss / n, "test_metric": correct / n}
        for k, v in out.items():
            logger.log_metric(self._self_addr, k, v)
        return out
