"""Learner contract and factory.

Reference: p2pfl/learning/frameworks/learner.py:33-167 and learner_factory.py:29-56.
``set_model`` differs on purpose: the reference *replaces* the learner's model object with the
aggregated one (rebuilding optimizer state around a deep copy); here the aggregated parameters are
*copied into the live slab* so the module, its gradient slab and any captured CUDA graph stay valid.
"""

from __future__ import annotations

import abc
from typing import Dict, List, Union

import numpy as np

from myfyp_b200.learning.callbacks.callback import CallbackFactory, P2PFLCallback
from myfyp_b200.learning.p2pfl_model import P2PFLModel


class Learner(abc.ABC):
    def __init__(self, model: P2PFLModel, data, self_addr: str = "unknown-node", aggregator=None) -> None:
        self.model: P2PFLModel = model
        self.data = data
        self._self_addr = self_addr
        self.callbacks: List[P2PFLCallback] = []
        if aggregator is not None:
            self.callbacks = CallbackFactory.create_callbacks(framework=self.get_framework(), aggregator=aggregator)
        self.epochs: int = 1

    def set_addr(self, addr: str) -> None:
        self._self_addr = addr

    def set_model(self, model: Union[P2PFLModel, List[np.ndarray], bytes, object]) -> None:
        if isinstance(model, P2PFLModel):
            if self.model is None:
                self.model = model
            else:
                self.model.set_parameters(model if hasattr(model, "slab") else model.get_parameters())
                self.model.additional_info.update(model.get_info())
                if model.contributors:
                    self.model.set_contribution(model.contributors, model.num_samples)
        else:
            self.model.set_parameters(model)
        self.update_callbacks_with_model_info()

    def get_model(self) -> P2PFLModel:
        return self.model

    def set_data(self, data) -> None:
        self.data = data

    def get_data(self):
        return self.data

    def set_epochs(self, epochs: int) -> None:
        self.epochs = epochs

    def update_callbacks_with_model_info(self) -> None:
        info = self.model.get_info()
        for cb in self.callbacks:
            if cb.get_name() in info:
                cb.set_info(info[cb.get_name()])

    def add_callback_info_to_model(self) -> None:
        for cb in self.callbacks:
            self.model.add_info(cb.get_name(), cb.get_info())

    @abc.abstractmethod
    def fit(self) -> P2PFLModel: ...

    @abc.abstractmethod
    def interrupt_fit(self) -> None: ...

    @abc.abstractmethod
    def evaluate(self) -> Dict[str, float]: ...

    @abc.abstractmethod
    def get_framework(self) -> str: ...


class LearnerFactory:
    """``model.get_framework()`` -> learner class (only the PyTorch binding exists)."""

    @staticmethod
    def create_learner(model: P2PFLModel):
        from myfyp_b200.learning.frameworks import learner_for

        return learner_for(model.get_framework())
