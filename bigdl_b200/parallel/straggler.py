"""Straggler mitigation math of the reference (S/optim/DistriOptimizer.scala:241-244,278,342-344,423-449,
S/utils/Util.scala:32 ``kthLargest``): after ``warmupIteration`` iterations the slowest ``dropPercentage`` of replica
computations are abandoned by a timeout equal to the k-th largest observed time; an iteration in which more than
``maxDropPercentage`` were dropped is discarded entirely.

``Optimizer.setDropModuleProperty`` builds the policy; the training loop (optim/optimizer.py ``_apply_drop_policy``)
exchanges the ranks' forward+backward times every iteration, drops the gradients of ranks above the threshold and
discards over-dropped iterations. A replica is a rank here (the reference drops per-core model replicas)."""
from __future__ import annotations

import random


def kth_largest(values, k: int):
    """k-th largest element, 1-based (quick-select like Util.kthLargest; k==0 returns +inf: drop nothing)."""
    if k <= 0:
        return float("inf")
    arr = list(values)
    if k > len(arr):
        raise ValueError("k exceeds the number of values")
    lo, hi, target = 0, len(arr) - 1, k - 1
    while True:
        if lo == hi:
            return arr[lo]
        p = arr[random.randint(lo, hi)]
        i, j = lo, hi
        while i <= j:
            while arr[i] > p:
                i += 1
            while arr[j] < p:
                j -= 1
            if i <= j:
                arr[i], arr[j] = arr[j], arr[i]
                i, j = i + 1, j - 1
        if target <= j:
            hi = j
        elif target >= i:
            lo = i
        else:
            return arr[target]


class DropPolicy:
    def __init__(self, drop_percentage: float, max_drop_percentage: float, batch_size: int = 100,
                 warmup_iteration: int = 200):
        if not 0 <= drop_percentage <= max_drop_percentage <= 1:
            raise ValueError("need 0 <= dropPercentage <= maxDropPercentage <= 1")
        self.drop, self.max_drop, self.window, self.warmup = drop_percentage, max_drop_percentage, batch_size, warmup_iteration
        self.times: list[float] = []
        self.threshold = float("inf")
        self.dropped_total = 0

    def record(self, iteration: int, replica_times) -> None:
        """Feed the per-replica compute times of one iteration; recompute the timeout every ``window`` iterations."""
        if iteration <= self.warmup:
            return
        self.times.extend(replica_times)
        per_iter = max(1, len(replica_times))
        if len(self.times) >= self.window * per_iter:
            k = int(self.drop * len(self.times))
            self.threshold = kth_largest(self.times, k) if k > 0 else float("inf")
            self.times = []

    def apply(self, replica_times):
        """→ (kept indices, discard_iteration)."""
        keep = [i for i, t in enumerate(replica_times) if t <= self.threshold]
        dropped = len(replica_times) - len(keep)
        self.dropped_total += dropped
        discard = dropped > self.max_drop * len(replica_times)
        return keep, discard
