"""Stateful model collector + the ``aggregate`` contract.

Reference: p2pfl/learning/aggregators/aggregator.py:35-270.  Semantics kept (SURVEY
App. C.1): a model is accepted iff the aggregation is still incomplete, *all* its
contributors are in the train set and *none* is already held; the finish event is set
when the held contributors cover the train set; ``get_model(except_nodes)`` offers a
partial aggregate (``partial_aggregation=True``) or one raw held model.

Fixes: empty contributor list returns ``[]`` instead of releasing an un-acquired lock
(Q4); ``wait_and_get_aggregation`` reads ``Settings.AGGREGATION_TIMEOUT`` at call time (Q5).
"""

from __future__ import annotations

import threading
from typing import Callable, List, Optional

from myfyp_b200.learning.p2pfl_model import P2PFLModel
from myfyp_b200.management.logger import logger
from myfyp_b200.settings import Settings


class NoModelsToAggregateError(Exception):
    """There is nothing (left) to aggregate."""


class Aggregator:
    def __init__(self, node_name: str = "unknown") -> None:
        self.node_name = node_name
        self.federation_size: Optional[int] = None  # nodes this node knows of (itself included); the train stage refreshes it every round
        self.__train_set: List[str] = []
        self.__models: List[P2PFLModel] = []
        self.partial_aggregation = False
        self.__agg_lock = threading.Lock()
        self._finish_aggregation_event = threading.Event()
        self._finish_aggregation_event.set()
        self._listeners: List[Callable[[], None]] = []

    # -- to implement ------------------------------------------------------------
    def aggregate(self, models: List[P2PFLModel]) -> P2PFLModel:
        raise NotImplementedError

    def get_required_callbacks(self) -> List[str]:
        return []

    # -- lifecycle -----------------------------------------------------------------
    def set_nodes_to_aggregate(self, nodes_to_aggregate: List[str]) -> None:
        if not self._finish_aggregation_event.is_set():
            raise Exception("It is not possible to set nodes to aggregate when the aggregation is running.")
        self.__train_set = list(nodes_to_aggregate)
        self._finish_aggregation_event.clear()

    def local_device(self, models: List[P2PFLModel]):
        """Where this node's aggregation should run: the GPU of the model this node contributed (its own snapshot), so that
        the result is born where the learner will adopt it; else the first model's device."""
        for m in models:
            try:
                if self.node_name in m.get_contributors() and hasattr(m, "slab"):
                    return m.slab.device
            except Exception:  # noqa: BLE001
                pass
        return getattr(getattr(models[0], "slab", None), "device", None) if models else None

    def observe_global(self, model: P2PFLModel) -> None:
        """The node adopted an aggregate it did not compute itself (it was not in the round's train set).  Aggregators with
        server-side state that must stay equal on all nodes (Scaffold's control variate) refresh it from the model's info."""

    def is_collecting(self) -> bool:
        """True between :meth:`set_nodes_to_aggregate` and the moment the last model of the train set is filed."""
        return not self._finish_aggregation_event.is_set()

    def clear(self) -> None:
        with self.__agg_lock:
            self.__train_set = []
            self.__models = []
            self._finish_aggregation_event.set()

    def on_model_added(self, fn: Callable[[], None]) -> None:
        self._listeners.append(fn)

    def drop_node(self, addr: str) -> bool:
        """A train-set member is known to be dead (evicted by the failure detector): stop waiting for its model.

        The reference only learns this by sitting out ``AGGREGATION_TIMEOUT`` (aggregator.py:192-208).  A model the node
        already delivered (alone or inside a partial aggregate) stays in; returns True if the train set shrank."""
        with self.__agg_lock:
            if addr not in self.__train_set or addr in self.get_aggregated_models():
                return False
            self.__train_set = [n for n in self.__train_set if n != addr]
            if self.__train_set and len(self.get_aggregated_models()) >= len(self.__train_set):
                self._finish_aggregation_event.set()
        for fn in self._listeners:
            fn()
        return True

    # -- collection ------------------------------------------------------------------
    def get_aggregated_models(self) -> List[str]:
        out: List[str] = []
        for m in list(self.__models):
            out += m.get_contributors()
        return out

    def _held(self) -> List[P2PFLModel]:
        return list(self.__models)

    def add_model(self, model: P2PFLModel) -> List[str]:
        try:
            contributors = model.get_contributors()
        except ValueError:
            contributors = []
        if not contributors:
            logger.debug(self.node_name, "Received a model without a list of contributors.")
            return []
        with self.__agg_lock:
            held = self.get_aggregated_models()
            if len(self.__train_set) <= len(held):
                logger.debug(self.node_name, "🚫 Received a model when is not needed (already aggregated).")
                return []
            if not all(n in self.__train_set for n in contributors):
                logger.debug(self.node_name, f"Can't add a model from a node ({contributors}) that is not in the training set.")
                return []
            if any(n in held for n in contributors):
                logger.debug(self.node_name, f"Can't add a model from a node ({contributors}) that is already in the training set.")
                return []
            self.__models.append(model)
            held = held + list(contributors)
            logger.info(self.node_name, f"🧩 Model added ({len(held)}/{len(self.__train_set)}) from {contributors}")
            if len(held) >= len(self.__train_set):
                self._finish_aggregation_event.set()
        for fn in self._listeners:
            fn()
        return held

    def wait_and_get_aggregation(self, timeout: Optional[float] = None) -> P2PFLModel:
        if timeout is None:
            timeout = Settings.AGGREGATION_TIMEOUT
        event_set = self._finish_aggregation_event.wait(timeout=timeout)
        missing = self.get_missing_models()
        if not event_set:
            logger.info(self.node_name, f"⏳ Aggregation wait timed out. Missing models: {missing}")
        elif missing:
            logger.info(self.node_name, f"❌ Aggregation event set, but missing models:  {missing}")
        else:
            logger.info(self.node_name, "🧠 Aggregating models.")
        return self.aggregate(self._held())

    def get_missing_models(self) -> set:
        return set(self.__train_set) - set(self.get_aggregated_models())

    def peek_offer(self, except_nodes: List[str]) -> List[str]:
        """Contributors of what :meth:`get_model` would return for a peer holding ``except_nodes`` - without aggregating."""
        usable = [m for m in self._held() if all(n not in except_nodes for n in m.get_contributors())]
        if not usable:
            return []
        if self.partial_aggregation:
            out: List[str] = []
            for m in usable:
                out += m.get_contributors()
            return out
        return list(usable[0].get_contributors())

    def get_model(self, except_nodes: List[str]) -> P2PFLModel:
        """What to offer a peer that already holds ``except_nodes``."""
        usable = [m for m in self._held() if all(n not in except_nodes for n in m.get_contributors())]
        if self.partial_aggregation:
            return self.aggregate(usable)
        if not usable:
            raise NoModelsToAggregateError("No remaining models available for aggregation.")
        return usable[0]
