"""Fixed worker pool with the reference's invocation verbs (S/utils/ThreadPool.scala:38-235): ``invokeAndWait`` (run
all, gather results), ``invokeAndWait2(tasks, timeout)`` (wait at most ``timeout`` seconds and report which tasks were
abandoned — the mechanism behind straggler dropping), ``invoke`` (fire and return futures) and ``sync``.
Host-side only: GPU work is ordered by CUDA streams, not by this pool; the C++ pool in csrc/host/thread_pool.cpp is its
native twin for the data-loader kernels."""
from __future__ import annotations

import concurrent.futures as cf
import os


class ThreadPool:
    def __init__(self, pool_size: int | None = None):
        self._size = pool_size or max(1, (os.cpu_count() or 2))
        self._ex = cf.ThreadPoolExecutor(max_workers=self._size, thread_name_prefix="bigdl-b200")
        self.mklThreads = 1

    def getPoolSize(self) -> int:
        return self._size

    def setPoolSize(self, size: int):
        self._ex.shutdown(wait=True)
        self._size = size
        self._ex = cf.ThreadPoolExecutor(max_workers=size, thread_name_prefix="bigdl-b200")
        return self

    def setMKLThread(self, n: int):
        """Per-worker math-library threads (MKL in the reference → torch intra-op threads here)."""
        import torch

        self.mklThreads = n
        torch.set_num_threads(max(1, n))
        return self

    def invoke(self, tasks):
        if callable(tasks):
            return self._ex.submit(tasks)
        return [self._ex.submit(t) for t in tasks]

    def invokeAndWait(self, tasks, timeout: float | None = None):
        futs = [self._ex.submit(t) for t in tasks]
        return [f.result(timeout=timeout) for f in futs]

    def invokeAndWait2(self, tasks, timeout: float | None = None):
        """Returns futures; tasks still running after ``timeout`` seconds are cancelled if not started and flagged
        ``f.abandoned = True`` either way (ThreadPool.scala:156-200)."""
        futs = [self._ex.submit(t) for t in tasks]
        done, not_done = cf.wait(futs, timeout=timeout)
        for f in futs:
            f.abandoned = f in not_done
            if f.abandoned:
                f.cancel()
        return futs

    def sync(self, futures, timeout: float | None = None):
        for f in futures:
            f.result(timeout=timeout)

    def shutdown(self):
        self._ex.shutdown(wait=False)
