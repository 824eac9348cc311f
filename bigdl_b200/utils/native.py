"""ctypes bindings of the C++ host runtime (include/bigdl_b200_host.h, csrc/host/*.cpp): thread pool, CRC32/CRC32C,
TFRecord framing, truncated-bf16 codec, MT19937 stream, multi-threaded image batcher, BigQuant int8 reference GEMM."""
from __future__ import annotations

import ctypes as C

import numpy as np

from ..ops import _lib

_sig_done = False


def lib():
    global _sig_done
    h = _lib.host_lib()
    if not _sig_done:
        P, I, Z, U32, U64, D, F = C.c_void_p, C.c_int, C.c_size_t, C.c_uint32, C.c_uint64, C.c_double, C.c_float
        sigs = {
            "bdlh_pool_size": ([], I), "bdlh_pool_resize": ([I], None),
            "bdlh_crc32c": ([P, Z, U32], U32), "bdlh_crc32c_masked": ([P, Z], U32), "bdlh_crc32": ([P, Z, U32], U32),
            "bdlh_tfrecord_frame": ([P, Z, P], Z), "bdlh_tfrecord_next": ([P, Z, Z, C.POINTER(Z), I], C.c_longlong),
            "bdlh_bf16_compress": ([P, P, Z], None), "bdlh_bf16_decompress": ([P, P, Z], None),
            "bdlh_bf16_add": ([P, P, Z], None),
            "bdlh_rng_new": ([U64], P), "bdlh_rng_free": ([P], None), "bdlh_rng_seed": ([P, U64], None),
            "bdlh_rng_u32": ([P], U32), "bdlh_rng_uniform": ([P, D, D], D), "bdlh_rng_normal": ([P, D, D], D),
            "bdlh_rng_fill_uniform": ([P, P, Z, D, D], None), "bdlh_rng_fill_normal": ([P, P, Z, D, D], None),
            "bdlh_rng_permutation": ([P, P, Z], None),
            "bdlh_batch_images": ([P, P, P, I, I, I, I, I, P, P, P, F, I, I, P], I),
            "bdlh_gather_rows": ([P, Z, I, P], None),
            "bdlh_quantize_rows_int8": ([P, P, P, I, I], None),
            "bdlh_int8_gemm": ([P, P, P, P, P, P, I, I, I], None),
            "bdlh_version": ([], C.c_char_p),
            "bdlh_set_affinity": ([P, I], I), "bdlh_get_affinity": ([P, I], I),
            "bdlh_logical_cores": ([], I), "bdlh_physical_cores": ([], I),
        }
        for name, (args, res) in sigs.items():
            fn = getattr(h, name)
            fn.argtypes, fn.restype = args, res
        _sig_done = True
    return h


def _buf(b):
    a = np.frombuffer(b, dtype=np.uint8) if not isinstance(b, np.ndarray) else b
    return a.ctypes.data if a.size else None, a.size, a


def crc32c(data: bytes, seed: int = 0) -> int:
    p, n, keep = _buf(data)
    return int(lib().bdlh_crc32c(p, n, seed))


def crc32c_masked(data: bytes) -> int:
    p, n, keep = _buf(data)
    return int(lib().bdlh_crc32c_masked(p, n))


def crc32(data: bytes, seed: int = 0) -> int:
    p, n, keep = _buf(data)
    return int(lib().bdlh_crc32(p, n, seed))


def tfrecord_frame(payload: bytes) -> bytes:
    out = np.empty(len(payload) + 16, dtype=np.uint8)
    p, n, keep = _buf(payload)
    lib().bdlh_tfrecord_frame(p, n, out.ctypes.data)
    return out.tobytes()


def tfrecord_iter(buf: bytes, verify: bool = True):
    """Yield the payloads of a TFRecord byte string; raises ValueError on a corrupt record."""
    arr = np.frombuffer(buf, dtype=np.uint8)
    pos, off = 0, C.c_size_t(0)
    h = lib()
    while True:
        n = h.bdlh_tfrecord_next(arr.ctypes.data if arr.size else None, arr.size, pos, C.byref(off), int(verify))
        if n == -1:
            return
        if n < 0:
            raise ValueError(f"corrupt TFRecord at byte {pos}")
        yield buf[off.value:off.value + n]
        pos = off.value + n + 4


def bf16_compress(x: np.ndarray) -> np.ndarray:
    x = np.ascontiguousarray(x, dtype=np.float32)
    out = np.empty(x.shape, dtype=np.uint16)
    lib().bdlh_bf16_compress(x.ctypes.data, out.ctypes.data, x.size)
    return out


def bf16_decompress(h: np.ndarray) -> np.ndarray:
    h = np.ascontiguousarray(h, dtype=np.uint16)
    out = np.empty(h.shape, dtype=np.float32)
    lib().bdlh_bf16_decompress(h.ctypes.data, out.ctypes.data, h.size)
    return out


def bf16_add_(acc: np.ndarray, other: np.ndarray) -> np.ndarray:
    assert acc.dtype == np.uint16 and other.dtype == np.uint16 and acc.flags.c_contiguous and other.flags.c_contiguous
    lib().bdlh_bf16_add(acc.ctypes.data, other.ctypes.data, acc.size)
    return acc


class NativeRNG:
    """MT19937 stream living in C++ (one per data-loader thread; the Python RNG in utils/random_generator.py keeps the
    global Torch-compatible stream)."""

    def __init__(self, seed: int = 1):
        self._h = lib().bdlh_rng_new(seed)

    def __del__(self):
        try:
            lib().bdlh_rng_free(self._h)
        except Exception:
            pass

    def seed(self, s: int):
        lib().bdlh_rng_seed(self._h, s)
        return self

    def random(self) -> int:
        return int(lib().bdlh_rng_u32(self._h))

    def uniform(self, a=0.0, b=1.0) -> float:
        return float(lib().bdlh_rng_uniform(self._h, a, b))

    def normal(self, mean=0.0, stdv=1.0) -> float:
        return float(lib().bdlh_rng_normal(self._h, mean, stdv))

    def fill_uniform(self, n: int, a=0.0, b=1.0) -> np.ndarray:
        out = np.empty(n, dtype=np.float32)
        lib().bdlh_rng_fill_uniform(self._h, out.ctypes.data, n, a, b)
        return out

    def fill_normal(self, n: int, mean=0.0, stdv=1.0) -> np.ndarray:
        out = np.empty(n, dtype=np.float32)
        lib().bdlh_rng_fill_normal(self._h, out.ctypes.data, n, mean, stdv)
        return out

    def permutation(self, n: int) -> np.ndarray:
        out = np.empty(n, dtype=np.int64)
        lib().bdlh_rng_permutation(self._h, out.ctypes.data, n)
        return out


def batch_images(images, crop_h: int, crop_w: int, mean, std, scale: float = 1.0, random_crop: bool = False,
                 random_flip: bool = False, seeds=None, nhwc: bool = False, out=None, bf16: bool = False):
    """images: list of uint8 HxWx3 arrays → [n,3,crop_h,crop_w] (or NHWC) float32 / bf16-bits(uint16) array.
    ``out`` may be a preallocated (pinned) buffer: a numpy array or a torch tensor of the right size."""
    import torch

    n = len(images)
    imgs = [np.ascontiguousarray(im, dtype=np.uint8) for im in images]
    ptrs = (C.c_void_p * n)(*[im.ctypes.data for im in imgs])
    hs = np.array([im.shape[0] for im in imgs], dtype=np.int32)
    ws = np.array([im.shape[1] for im in imgs], dtype=np.int32)
    shape = (n, crop_h, crop_w, 3) if nhwc else (n, 3, crop_h, crop_w)
    if out is None:
        out = np.empty(shape, dtype=np.uint16 if bf16 else np.float32)
    optr = out.data_ptr() if isinstance(out, torch.Tensor) else out.ctypes.data
    mean = np.ascontiguousarray(mean, dtype=np.float32)
    std = np.ascontiguousarray(std, dtype=np.float32)
    sd = None if seeds is None else np.ascontiguousarray(seeds, dtype=np.uint64)
    rc = lib().bdlh_batch_images(C.cast(ptrs, C.c_void_p), hs.ctypes.data, ws.ctypes.data, n, crop_h, crop_w,
                                 int(random_crop), int(random_flip), None if sd is None else sd.ctypes.data,
                                 mean.ctypes.data, std.ctypes.data, float(scale), int(nhwc), int(bf16), optr)
    if rc != 0:
        raise ValueError("an image is smaller than the crop size")
    return out


def quantize_rows_int8(x: np.ndarray):
    x = np.ascontiguousarray(x, dtype=np.float32)
    q = np.empty(x.shape, dtype=np.int8)
    am = np.empty(x.shape[0], dtype=np.float32)
    lib().bdlh_quantize_rows_int8(x.ctypes.data, q.ctypes.data, am.ctypes.data, x.shape[0], x.shape[1])
    return q, am


def int8_gemm(a_q, a_absmax, b_q, b_absmax, bias=None) -> np.ndarray:
    M, K = a_q.shape
    N = b_q.shape[0]
    out = np.empty((M, N), dtype=np.float32)
    b = None if bias is None else np.ascontiguousarray(bias, dtype=np.float32)
    lib().bdlh_int8_gemm(a_q.ctypes.data, a_absmax.ctypes.data, b_q.ctypes.data, b_absmax.ctypes.data,
                         None if b is None else b.ctypes.data, out.ctypes.data, M, N, K)
    return out


def pool_size() -> int:
    return int(lib().bdlh_pool_size())


def pool_resize(n: int):
    lib().bdlh_pool_resize(int(n))


def set_affinity(cpus) -> None:
    """Pin the calling thread to the given logical CPUs (reference: MklDnn.setAffinity, affinity.c)."""
    arr = (C.c_int * len(cpus))(*[int(c) for c in cpus])
    if lib().bdlh_set_affinity(arr, len(cpus)) != 0:
        raise OSError("sched_setaffinity failed")


def get_affinity() -> list[int]:
    arr = (C.c_int * 4096)()
    n = lib().bdlh_get_affinity(arr, 4096)
    if n < 0:
        raise OSError("sched_getaffinity failed")
    return [arr[i] for i in range(min(n, 4096))]


def logical_cores() -> int:
    return int(lib().bdlh_logical_cores())


def physical_cores() -> int:
    """Distinct (package, core) pairs — what the reference's Engine uses for its default core number."""
    return int(lib().bdlh_physical_cores())
