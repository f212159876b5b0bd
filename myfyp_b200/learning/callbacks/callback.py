"""Callback contract + registry.

Reference: learning/frameworks/callback.py:25-50 and callback_factory.py:32-101 - an
aggregator names the callbacks it needs (``get_required_callbacks``) and the learner gets
them from the factory.  Callbacks exchange state with the aggregator through the model's
``additional_info[<callback name>]``.

B200-first: callbacks do not run Python per batch.  They contribute *terms to the fused
optimizer kernel* (``optimizer_terms``) and run only at fit start / fit end.
"""

from __future__ import annotations

from typing import Any, Callable, Dict, List, Type


class P2PFLCallback:
    def __init__(self) -> None:
        self.additional_info: Dict[str, Any] = {}

    @staticmethod
    def get_name() -> str:
        raise NotImplementedError

    def get_info(self) -> Any:
        return self.additional_info

    def set_info(self, info: Any) -> None:
        self.additional_info = info

    # -- hooks used by the native training loop ---------------------------------------
    def on_fit_start(self, learner) -> None: ...

    def optimizer_terms(self, learner) -> Dict[str, Any]:
        """Extra operands for the fused optimizer step (e.g. prox_mu / prox_ref / cv_global / cv_local)."""
        return {}

    def on_fit_end(self, learner, steps: int, lr: float) -> None: ...


class CallbackFactory:
    """``framework -> name -> constructor`` registry."""

    _registry: Dict[str, Dict[str, Callable[..., P2PFLCallback]]] = {}

    @classmethod
    def register_callback(cls, learner: str, callback: Type[P2PFLCallback]) -> None:
        cls._registry.setdefault(learner, {})[callback.get_name()] = callback

    @classmethod
    def create_callbacks(cls, framework: str, aggregator) -> List[P2PFLCallback]:
        out: List[P2PFLCallback] = []
        for name in aggregator.get_required_callbacks():
            try:
                ctor = cls._registry[framework][name]
            except KeyError as e:
                raise ValueError(f"Callback {name} is not registered for framework {framework}.") from e
            try:
                out.append(ctor(aggregator=aggregator))
            except TypeError:
                out.append(ctor())
        return out
