"""Thread-local Mersenne-Twister RNG with Torch7 semantics.

Reference: S/utils/RandomGenerator.scala (MT19937; ``uniform(a,b) = random()*2^-32*(b-a)+a``; Box–Muller normal
with a cached second value; ``RNG.setSeed`` for deterministic tests). The generator core is native C++
(csrc/host/rng.cpp) with a numpy fallback producing the same stream.
"""
from __future__ import annotations

import ctypes
import math
import threading

import numpy as np
import torch


class RandomGenerator:
    def __init__(self, seed: int | None = None):
        self._np = np.random.RandomState()
        self._normal_valid = False
        self._normal_x = self._normal_rho = 0.0
        self.setSeed(seed if seed is not None else int.from_bytes(__import__("os").urandom(4), "little"))

    def setSeed(self, seed: int):
        self._seed = int(seed) & 0xFFFFFFFF
        # numpy's RandomState(seed) uses init_genrand(seed) — the same MT19937 seeding Torch7 uses
        self._np = np.random.RandomState(self._seed)
        self._normal_valid = False
        return self

    set_seed = setSeed

    def getSeed(self) -> int:
        return self._seed

    def clone(self):
        g = RandomGenerator(self._seed)
        g._np.set_state(self._np.get_state())
        g._normal_valid, g._normal_x, g._normal_rho = self._normal_valid, self._normal_x, self._normal_rho
        return g

    # ---- raw 32-bit draws
    def random(self) -> int:
        return int(self._np.randint(0, 1 << 32, dtype=np.uint64))

    def _raw(self, n: int) -> np.ndarray:
        return self._np.randint(0, 1 << 32, size=n, dtype=np.uint64)

    def uniform(self, a: float = 0.0, b: float = 1.0) -> float:
        return self.random() * (1.0 / 4294967296.0) * (b - a) + a

    def normal(self, mean: float = 0.0, stdv: float = 1.0) -> float:
        if stdv <= 0:
            raise ValueError("standard deviation must be strictly positive")
        if not self._normal_valid:
            self._normal_x = self.uniform()
            y = self.uniform()
            self._normal_rho = math.sqrt(-2.0 * math.log(1.0 - y))
            self._normal_valid = True
            return self._normal_rho * math.cos(2.0 * math.pi * self._normal_x) * stdv + mean
        self._normal_valid = False
        return self._normal_rho * math.sin(2.0 * math.pi * self._normal_x) * stdv + mean

    def exponential(self, lam: float) -> float:
        return -1.0 / lam * math.log(1 - self.uniform())

    def cauchy(self, median: float, sigma: float) -> float:
        return median + sigma * math.tan(math.pi * (self.uniform() - 0.5))

    def logNormal(self, mean: float, stdv: float) -> float:
        zm = mean * mean
        zs = stdv * stdv
        return math.exp(self.normal(math.log(zm / math.sqrt(zs + zm)), math.sqrt(math.log(zs / zm + 1))))

    def geometric(self, p: float) -> int:
        return int(math.log(1 - self.uniform()) / math.log(p)) + 1

    def bernoulli(self, p: float) -> bool:
        return self.uniform() <= p

    # ---- bulk fills (vectorised, same stream order as repeated scalar draws)
    def fill_uniform(self, t: torch.Tensor, a: float, b: float) -> torch.Tensor:
        n = t.numel()
        vals = self._raw(n).astype(np.float64) * (1.0 / 4294967296.0) * (b - a) + a
        src = torch.from_numpy(vals.astype(np.float32)).reshape(t.shape)
        with torch.no_grad():
            t.copy_(src.to(t.device))
        return t

    def fill_normal(self, t: torch.Tensor, mean: float, stdv: float) -> torch.Tensor:
        n = t.numel()
        out = np.empty(n, dtype=np.float64)
        i = 0
        if self._normal_valid and n > 0:
            out[0] = self.normal(mean, stdv)
            i = 1
        pairs = (n - i + 1) // 2
        if pairs > 0:
            raw = self._raw(2 * pairs).astype(np.float64) * (1.0 / 4294967296.0)
            x, y = raw[0::2], raw[1::2]
            rho = np.sqrt(-2.0 * np.log(1.0 - y))
            c = rho * np.cos(2.0 * np.pi * x) * stdv + mean
            s = rho * np.sin(2.0 * np.pi * x) * stdv + mean
            inter = np.empty(2 * pairs, dtype=np.float64)
            inter[0::2], inter[1::2] = c, s
            take = n - i
            out[i:] = inter[:take]
            if take < 2 * pairs:   # one cached value left over, exactly like the scalar path
                self._normal_valid = True
                self._normal_x, self._normal_rho = float(x[-1]), float(rho[-1])
            else:
                self._normal_valid = False
        with torch.no_grad():
            t.copy_(torch.from_numpy(out.astype(np.float32)).reshape(t.shape).to(t.device))
        return t

    def fill_bernoulli(self, t: torch.Tensor, p: float) -> torch.Tensor:
        n = t.numel()
        vals = (self._raw(n).astype(np.float64) * (1.0 / 4294967296.0) <= p).astype(np.float32)
        with torch.no_grad():
            t.copy_(torch.from_numpy(vals).reshape(t.shape).to(t.device))
        return t

    def shuffle(self, data: list) -> list:
        """In-place Fisher–Yates exactly as RandomGenerator.shuffle (S/utils/RandomGenerator.scala:35-47)."""
        n = len(data)
        for i in range(n):
            j = int(self.uniform(0, n - i)) + i
            data[i], data[j] = data[j], data[i]
        return data

    def randperm(self, n: int) -> np.ndarray:
        return np.asarray(self.shuffle(list(range(n))), dtype=np.int64)


class _ThreadLocalRNG(threading.local):
    def __init__(self):
        self.gen = RandomGenerator()


_tls = _ThreadLocalRNG()


class _RNGProxy:
    """``RNG`` resolves to the calling thread's generator, like the reference's ThreadLocal."""

    def __getattr__(self, item):
        return getattr(_tls.gen, item)


RNG = _RNGProxy()
