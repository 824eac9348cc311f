"""Triggers over the optimizer state table (keys ``epoch`` (starts at 1), ``neval`` (next iteration index, starts at
1), ``Loss``, ``score``). Reference: S/optim/Trigger.scala:30-130; Python names from P/optim/optimizer.py."""
from __future__ import annotations


class Trigger:
    def __call__(self, state) -> bool:  # pragma: no cover
        raise NotImplementedError

    # reference-style factories
    @staticmethod
    def everyEpoch():
        return EveryEpoch()

    @staticmethod
    def severalIteration(interval: int):
        return SeveralIteration(interval)

    @staticmethod
    def maxEpoch(n: int):
        return MaxEpoch(n)

    @staticmethod
    def maxIteration(n: int):
        return MaxIteration(n)

    @staticmethod
    def maxScore(s: float):
        return MaxScore(s)

    @staticmethod
    def minLoss(l: float):
        return MinLoss(l)

    @staticmethod
    def and_(first, *others):
        return TriggerAnd(first, *others)

    @staticmethod
    def or_(first, *others):
        return TriggerOr(first, *others)


class EveryEpoch(Trigger):
    """Fires once each time the epoch counter advances."""

    def __init__(self):
        self.last_epoch = -1

    def __call__(self, state):
        e = state["epoch"]
        if self.last_epoch == -1:
            self.last_epoch = e
            return False
        if e == self.last_epoch:
            return False
        self.last_epoch = e
        return True


class SeveralIteration(Trigger):
    def __init__(self, interval: int):
        if interval <= 0:
            raise ValueError("interval should be positive")
        self.interval = interval

    def __call__(self, state):
        n = state["neval"]
        return n != 0 and n % self.interval == 0


class MaxEpoch(Trigger):
    def __init__(self, max_epoch: int):
        self.max_epoch = max_epoch

    def __call__(self, state):
        return state["epoch"] > self.max_epoch


class MaxIteration(Trigger):
    def __init__(self, max_iteration: int):
        self.max_iteration = max_iteration

    def __call__(self, state):
        return state["neval"] > self.max_iteration


class MaxScore(Trigger):
    def __init__(self, max_score: float):
        self.max_score = max_score

    def __call__(self, state):
        return state.getOrElse("score", float("-inf")) > self.max_score


class MinLoss(Trigger):
    def __init__(self, min_loss: float):
        self.min_loss = min_loss

    def __call__(self, state):
        return state.getOrElse("Loss", float("inf")) < self.min_loss


class TriggerAnd(Trigger):
    def __init__(self, first, *others):
        self.triggers = [first, *others]

    def __call__(self, state):
        return all(t(state) for t in self.triggers)


class TriggerOr(Trigger):
    def __init__(self, first, *others):
        self.triggers = [first, *others]

    def __call__(self, state):
        return any(t(state) for t in self.triggers)
