"""Experiment bookkeeping (reference: p2pfl/experiment.py:21-74)."""

from __future__ import annotations


class Experiment:
    """Name + total rounds + current round counter."""

    def __init__(self, exp_name: str, total_rounds: int, round: int = 0) -> None:
        self.exp_name = exp_name
        self.total_rounds = total_rounds
        self.round = round

    def increase_round(self) -> None:
        if self.round is None:
            raise ValueError("Round not initialized")
        self.round += 1

    def __str__(self) -> str:
        return f"Experiment(exp_name={self.exp_name}, total_rounds={self.total_rounds}, round={self.round})"

    __repr__ = __str__
