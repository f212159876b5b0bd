"""Aggregators: FedAvg, FedMedian / TrimmedMean, Krum / Multi-Krum, Scaffold (+ FedProx = FedAvg + callback)."""

from myfyp_b200.learning.aggregators.aggregator import Aggregator, NoModelsToAggregateError
from myfyp_b200.learning.aggregators.fedavg import FedAvg
from myfyp_b200.learning.aggregators.fedmedian import FedMedian, TrimmedMean
from myfyp_b200.learning.aggregators.krum import Krum
from myfyp_b200.learning.aggregators.scaffold import Scaffold
from myfyp_b200.learning.aggregators.fedprox import FedProx

__all__ = ["Aggregator", "NoModelsToAggregateError", "FedAvg", "FedMedian", "TrimmedMean", "Krum", "Scaffold", "FedProx"]
