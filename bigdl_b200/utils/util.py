"""Assorted helpers (S/utils/Util.scala:32-268, HashFunc.scala, Crc32.scala, LayerException.scala)."""
from __future__ import annotations

import struct

from ..parallel.straggler import kth_largest as kthLargest  # noqa: F401  (Util.kthLargest)


def getAndClearWeightBias(model):
    """Detach the parameter tensors from a model so its structure can be shipped separately (Util.scala:75-140)."""
    w, _ = model.getParameters()
    flat = w.detach().clone()
    return flat


def putWeightBias(flat, model):
    w, _ = model.getParameters()
    w.copy_(flat)
    model._mark_weights_dirty()
    return model


def initGradWeightBias(model):
    model.zeroGradParameters()
    return model


def shift(values, from_: int, to: int):
    """Move element ``from_`` to index ``to`` (Util.shift)."""
    v = list(values)
    v.insert(to, v.pop(from_))
    return v


class HashFunc:
    @staticmethod
    def murmurhash3_32(data, seed: int = 0) -> int:
        """MurmurHash3 x86_32 (S/utils/HashFunc.scala) of bytes / str / int — used by the hash-bucket feature columns."""
        if isinstance(data, str):
            data = data.encode("utf-8")
        elif isinstance(data, int):
            data = struct.pack("<i", data & 0xFFFFFFFF if data >= 0 else data)
        c1, c2, h = 0xCC9E2D51, 0x1B873593, seed & 0xFFFFFFFF
        n = len(data)
        for i in range(0, n - n % 4, 4):
            k = struct.unpack_from("<I", data, i)[0]
            k = (k * c1) & 0xFFFFFFFF
            k = ((k << 15) | (k >> 17)) & 0xFFFFFFFF
            k = (k * c2) & 0xFFFFFFFF
            h ^= k
            h = ((h << 13) | (h >> 19)) & 0xFFFFFFFF
            h = (h * 5 + 0xE6546B64) & 0xFFFFFFFF
        tail, k = data[n - n % 4:], 0
        for i, b in enumerate(tail):
            k |= b << (8 * i)
        if tail:
            k = (k * c1) & 0xFFFFFFFF
            k = ((k << 15) | (k >> 17)) & 0xFFFFFFFF
            k = (k * c2) & 0xFFFFFFFF
            h ^= k
        h ^= n
        h ^= h >> 16
        h = (h * 0x85EBCA6B) & 0xFFFFFFFF
        h ^= h >> 13
        h = (h * 0xC2B2AE35) & 0xFFFFFFFF
        h ^= h >> 16
        return h


class Crc32:
    """CRC32C with TFRecord masking (S/utils/Crc32.scala → netty Crc32c in the reference)."""

    @staticmethod
    def maskedCRC32(data: bytes) -> int:
        from . import native

        return native.crc32c_masked(data)

    @staticmethod
    def crc32c(data: bytes) -> int:
        from . import native

        return native.crc32c(data)


class LayerException(RuntimeError):
    """Exception tagged with the path of the layer that raised it (S/utils/LayerException.scala)."""

    def __init__(self, layer_msg: str, error: BaseException):
        super().__init__(f"Layer info: {layer_msg}\n{type(error).__name__}: {error}")
        self.layerMsg, self.error = layer_msg, error
