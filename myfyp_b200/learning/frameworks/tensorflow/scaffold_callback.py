"""Keras side of SCAFFOLD (reference: tensorflow/callbacks/scaffold_callback.py:30-163): an optimizer wrapper that corrects every
gradient with the control variates, and the callback that installs it at ``on_train_begin`` and publishes ``delta_y_i`` /
``delta_c_i`` at ``on_train_end``.  The arithmetic lives in :mod:`myfyp_b200.learning.frameworks.scaffold_math` (tested without
TensorFlow); this module is the thin Keras adaptor and needs ``tensorflow``, which this image does not ship."""

from __future__ import annotations

from typing import Any, Dict, List, Optional

import tensorflow as tf

from myfyp_b200.learning.callbacks.callback import CallbackFactory, P2PFLCallback
from myfyp_b200.learning.frameworks import scaffold_math


class ScaffoldOptimizerWrapper:
    """Stands in for the compiled optimizer: ``apply_gradients`` adds ``eta_l * (c_i - c)`` to every gradient first; everything else
    (``iterations``, ``learning_rate``, ``variables`` ...) is the wrapped optimizer's."""

    def __init__(self, optimizer, c_local: List[Any], c_global: List[Any], eta_l: float) -> None:
        self.__dict__.update(_inner=optimizer, _c_local=c_local, _c_global=c_global, _eta=float(eta_l))

    def apply_gradients(self, grads_and_vars, *args, **kwargs):
        pairs = list(grads_and_vars)
        fixed = scaffold_math.corrected_gradients([g for g, _ in pairs], self._c_local, self._c_global, self._eta)
        return self._inner.apply_gradients([(g, v) for g, (_, v) in zip(fixed, pairs)], *args, **kwargs)

    def __getattr__(self, name):
        return getattr(self.__dict__["_inner"], name)

    def __setattr__(self, name, value):
        setattr(self.__dict__["_inner"], name, value)


class SCAFFOLDCallback(tf.keras.callbacks.Callback, P2PFLCallback):
    def __init__(self, aggregator=None) -> None:
        tf.keras.callbacks.Callback.__init__(self)
        self.additional_info: Dict[str, Any] = {}
        self.c_local: List[tf.Variable] = []
        self.c_global: List[tf.Variable] = []
        self._x: List[Any] = []
        self._steps = 0
        self._eta: Optional[float] = None
        self._plain_optimizer = None

    @staticmethod
    def get_name() -> str:
        return "scaffold"

    def set_additional_info(self, info: Dict[str, Any]) -> None:
        self.additional_info = info

    def on_train_begin(self, logs=None) -> None:
        opt = self.model.optimizer
        opt = getattr(opt, "_inner", opt)  # a previous fit left our wrapper in place
        lr = opt.learning_rate
        self._eta = float(tf.keras.backend.get_value(lr(opt.iterations) if callable(lr) else lr))
        tv = self.model.trainable_variables
        if len(self.c_local) != len(tv):
            self.c_local = [tf.Variable(tf.zeros_like(v), trainable=False) for v in tv]
        g = self.additional_info.get("global_c") if isinstance(self.additional_info, dict) else None
        if g is not None:
            self.c_global = [tf.Variable(tf.convert_to_tensor(a, dtype=v.dtype), trainable=False) for a, v in zip(g, tv)]
        elif len(self.c_global) != len(tv):
            self.c_global = [tf.Variable(tf.zeros_like(v), trainable=False) for v in tv]
        self._x = [v.numpy().copy() for v in tv]
        self._steps = 0
        self._plain_optimizer = opt
        self.model.optimizer = ScaffoldOptimizerWrapper(opt, self.c_local, self.c_global, self._eta)

    def on_train_batch_end(self, batch, logs=None) -> None:
        self._steps += 1

    def on_train_end(self, logs=None) -> None:
        tv = self.model.trainable_variables
        new_c, dy, dc = scaffold_math.finalize(self._x, [v.numpy() for v in tv], [c.numpy() for c in self.c_local], max(self._steps, 1), self._eta)
        for var, val in zip(self.c_local, new_c):
            var.assign(val)
        if self._plain_optimizer is not None:
            self.model.optimizer = self._plain_optimizer
        self.additional_info = {"delta_y_i": dy, "delta_c_i": dc}


CallbackFactory.register_callback("tensorflow", SCAFFOLDCallback)
