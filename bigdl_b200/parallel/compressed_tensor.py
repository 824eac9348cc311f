"""Gradient/weight wire format: truncated bfloat16 ("fp16" in the reference's naming).

Reference: S/parameters/Parameter.scala:30-74 (``CompressedTensor`` trait, ``SerializerInstance`` with only "fp16"
registered), S/parameters/FP16CompressedTensor.scala:43-279 (compress/deCompress/add/parAdd, keep the upper 16 bits of
each float, sums re-truncate after every add), S/parameters/FP16SplitsCompressedTensor.scala (K split arrays; ``add`` is
not implemented there either). GPU tensors use the CUDA codec (csrc/elementwise.cu), CPU tensors the threaded C++ codec
(csrc/host/bf16_codec.cpp)."""
from __future__ import annotations

import numpy as np
import torch


def _trunc_bits(x: torch.Tensor) -> torch.Tensor:
    return (x.contiguous().view(torch.int32) >> 16).to(torch.int16)


def _expand_bits(b: torch.Tensor) -> torch.Tensor:
    return (b.to(torch.int32) << 16).view(torch.float32)


class CompressedTensor:
    def bytes(self, offset: int = 0, length: int | None = None) -> bytes:
        raise NotImplementedError

    def compress(self, src: torch.Tensor, offset: int = 0, length: int | None = None):
        raise NotImplementedError

    def deCompress(self, dst: torch.Tensor, src_offset: int = 0, length: int | None = None):
        raise NotImplementedError

    def add(self, other, offset: int = 0, length: int | None = None):
        raise NotImplementedError

    def parAdd(self, other, offset: int = 0, length: int | None = None):
        return self.add(other, offset, length)


class FP16CompressedTensor(CompressedTensor):
    """``length`` bf16 values held as int16 bit patterns (2 bytes each, big-endian when exported via ``bytes``)."""

    def __init__(self, source):
        if isinstance(source, int):
            self.buf = torch.zeros(source, dtype=torch.int16)
        elif isinstance(source, (bytes, bytearray)):
            be = np.frombuffer(source, dtype=">i2").astype(np.int16)
            self.buf = torch.from_numpy(be.copy())
        elif isinstance(source, torch.Tensor) and source.dtype == torch.int16:
            self.buf = source
        elif isinstance(source, torch.Tensor):
            self.buf = torch.zeros(source.numel(), dtype=torch.int16, device=source.device)
            self.compress(source)
        else:
            raise TypeError(type(source))

    def __len__(self):
        return self.buf.numel()

    def compress(self, src: torch.Tensor, offset: int = 0, length: int | None = None):
        flat = src.detach().reshape(-1).float()
        n = flat.numel() if length is None else length
        assert offset + n <= self.buf.numel()
        if flat.is_cuda:
            from ..ops import elementwise as ew

            out = self.buf[offset:offset + n].view(torch.bfloat16)
            ew.cast_f32_to_bf16(flat[:n].contiguous(), out, True)
        elif n >= 1 << 16:
            from ..utils import native

            self.buf[offset:offset + n] = torch.from_numpy(
                native.bf16_compress(flat[:n].contiguous().numpy()).view(np.int16))
        else:
            self.buf[offset:offset + n] = _trunc_bits(flat[:n])
        return self

    def deCompress(self, dst: torch.Tensor, src_offset: int = 0, length: int | None = None):
        n = dst.numel() if length is None else length
        out = _expand_bits(self.buf[src_offset:src_offset + n])
        dst.reshape(-1)[:n].copy_(out)
        return dst

    def add(self, other, offset: int = 0, length: int | None = None):
        ob = other.buf if isinstance(other, FP16CompressedTensor) else FP16CompressedTensor(other).buf
        n = ob.numel() if length is None else length
        a = self.buf[offset:offset + n]
        if not a.is_cuda and n >= 1 << 16 and a.is_contiguous() and ob.is_contiguous():
            from ..utils import native

            native.bf16_add_(a.numpy().view(np.uint16), ob[:n].numpy().view(np.uint16))
        else:
            a.copy_(_trunc_bits(_expand_bits(a) + _expand_bits(ob[:n])))
        return self

    def bytes(self, offset: int = 0, length: int | None = None) -> bytes:
        n = self.buf.numel() - offset if length is None else length
        return self.buf[offset:offset + n].cpu().numpy().astype(">i2").tobytes()


class FP16SplitsCompressedTensor(CompressedTensor):
    """The same payload stored as ``splits`` arrays (FP16SplitsCompressedTensor.scala:24-148); ``add`` is unsupported
    exactly as in the reference."""

    def __init__(self, length: int, splits: int):
        self.length, self.splits = length, splits
        per = (length + splits - 1) // splits
        self.parts = [FP16CompressedTensor(min(per, max(0, length - i * per))) for i in range(splits)]
        self._per = per

    def compress(self, src: torch.Tensor, offset: int = 0, length: int | None = None):
        flat = src.detach().reshape(-1)
        n = flat.numel() if length is None else length
        pos = 0
        while pos < n:
            gi = offset + pos
            p, po = divmod(gi, self._per)
            take = min(self._per - po, n - pos)
            self.parts[p].compress(flat[pos:pos + take], po, take)
            pos += take
        return self

    def deCompress(self, dst: torch.Tensor, src_offset: int = 0, length: int | None = None):
        flat = dst.reshape(-1)
        n = flat.numel() if length is None else length
        pos = 0
        while pos < n:
            gi = src_offset + pos
            p, po = divmod(gi, self._per)
            take = min(self._per - po, n - pos)
            self.parts[p].deCompress(flat[pos:pos + take], po, take)
            pos += take
        return dst

    def add(self, other, offset: int = 0, length: int | None = None):
        raise NotImplementedError("FP16SplitsCompressedTensor.add is not implemented (as in the reference)")

    def bytes(self, offset: int = 0, length: int | None = None) -> bytes:
        whole = b"".join(p.bytes() for p in self.parts)
        n = self.length - offset if length is None else length
        return whole[offset * 2:(offset + n) * 2]


class SerializerInstance:
    """Factory keyed by codec name (Parameter.scala:48-74); only "fp16" exists."""

    @staticmethod
    def create(source, pm: str = "fp16", splits: int | None = None):
        if pm != "fp16":
            raise ValueError("Unsupported parameter type")
        if splits is not None:
            assert isinstance(source, int)
            return FP16SplitsCompressedTensor(source, splits)
        return FP16CompressedTensor(source)
