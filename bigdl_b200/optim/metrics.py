"""Named timers / counters (S/optim/Metrics.scala:31-123). ``local`` metrics are plain floats; ``distributed`` ones
are all-reduced across ranks on ``summary()`` (the Spark-accumulator role)."""
from __future__ import annotations

import threading


class Metrics:
    def __init__(self):
        self._local: dict[str, list[float]] = {}
        self._dist: dict[str, list[float]] = {}
        self._lock = threading.Lock()

    def set(self, name: str, value: float, parallel: int = 1, distributed: bool = False):
        with self._lock:
            (self._dist if distributed else self._local)[name] = [float(value), float(parallel)]
        return self

    def add(self, name: str, value: float):
        with self._lock:
            if name in self._local:
                self._local[name][0] += value
            elif name in self._dist:
                self._dist[name][0] += value
            else:
                self._local[name] = [float(value), 1.0]
        return self

    def get(self, name: str):
        d = self._local.get(name) or self._dist.get(name)
        if d is None:
            raise KeyError(name)
        return d[0], int(d[1])

    def summary(self, unit: str = "s", scale: float = 1e9) -> str:
        import torch.distributed as dist

        lines = ["========== Metrics Summary =========="]
        for k, (v, p) in self._local.items():
            lines.append(f"{k} : {v / p / scale} {unit}")
        for k, (v, p) in self._dist.items():
            if dist.is_available() and dist.is_initialized():
                import torch

                t = torch.tensor([v], dtype=torch.float64)
                if dist.get_backend() == "nccl":
                    t = t.cuda()
                dist.all_reduce(t)
                v = float(t.item())
                p = p * dist.get_world_size()
            lines.append(f"{k} : {v / p / scale} {unit}")
        lines.append("=====================================")
        return "\n".join(lines)
