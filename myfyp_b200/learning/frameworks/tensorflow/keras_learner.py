"""``KerasLearner`` (reference: tensorflow/keras_learner.py:36-124): compile once, ``model.fit`` per round, metrics to the logger."""

from __future__ import annotations

from typing import Dict

import tensorflow as tf

from myfyp_b200.learning.frameworks.tensorflow.keras_dataset import KerasExportStrategy
from myfyp_b200.learning.learner import Learner
from myfyp_b200.management.logger import logger


class _FederatedLogger(tf.keras.callbacks.Callback):
    """Per-batch loss / metrics of ``model.fit`` -> the node's local metric store (reference: callbacks/keras_logger.py)."""

    def __init__(self, addr: str) -> None:
        super().__init__()
        self.addr, self.step = addr, 0

    def on_train_batch_end(self, batch, logs=None):
        for k, v in (logs or {}).items():
            logger.log_metric(self.addr, "train_" + k, float(v), step=self.step)
        self.step += 1


class KerasLearner(Learner):
    def __init__(self, model, data, self_addr: str = "unknown-node", aggregator=None, batch_size: int = 32, lr: float = 1e-3, **kwargs) -> None:
        super().__init__(model, data, self_addr, aggregator)
        self.batch_size = batch_size
        self._stop = False
        self.model.get_model().compile(
            optimizer=tf.keras.optimizers.Adam(lr), loss=tf.keras.losses.SparseCategoricalCrossentropy(from_logits=True), metrics=["sparse_categorical_accuracy"]
        )

    def get_framework(self) -> str:
        return "tensorflow"

    def fit(self):
        self._stop = False
        ds = self.data.export(KerasExportStrategy, train=True, batch_size=self.batch_size)
        stopper = tf.keras.callbacks.LambdaCallback(on_batch_end=lambda b, l: setattr(self.model.get_model(), "stop_training", self._stop))
        if self.epochs > 0:
            self.model.get_model().fit(ds, epochs=self.epochs, verbose=0, callbacks=[_FederatedLogger(self._self_addr), stopper, *self.callbacks])
        self.model.set_contribution([self._self_addr], self.data.get_num_samples(train=True))
        self.add_callback_info_to_model()
        return self.model

    def interrupt_fit(self) -> None:
        self._stop = True

    def evaluate(self) -> Dict[str, float]:
        ds = self.data.export(KerasExportStrategy, train=False, batch_size=max(self.batch_size, 256))
        loss, acc = self.model.get_model().evaluate(ds, verbose=0)
        out = {"test_loss": float(loss), "test_metric": float(acc)}
        for k, v in out.items():
            logger.log_metric(self._self_addr, k, v)
        return out
