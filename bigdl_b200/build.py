"""Ahead-of-time build of the native libraries, in-tree.

* ``bigdl_b200/lib/libbigdl_b200.so``       — sm_100a CUDA kernels (csrc/*.cu), plain C ABI (include/bigdl_b200.h)
* ``bigdl_b200/lib/libbigdl_b200_host.so``  — C++17 host runtime (csrc/host/*.cpp): RNG, bf16 codec, batcher,
  CRC32C/record IO, int8 reference GEMM. No CUDA dependency, so it loads on CPU-only machines.

The reference builds its natives with per-platform Makefiles under core/*/Makefile.common and CMake for MKL-DNN;
here a single nvcc/g++ invocation per translation unit is enough (nvcc cross-compiles without a GPU).
"""
from __future__ import annotations

import hashlib
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
CSRC = ROOT / "csrc"
INCLUDE = ROOT / "include"
LIBDIR = ROOT / "bigdl_b200" / "lib"
OBJDIR = ROOT / "build" / "obj"

NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
CXX = os.environ.get("CXX", "g++")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "--expt-relaxed-constexpr",
    "-Xcompiler", "-fPIC",
    f"-I{INCLUDE}", f"-I{CSRC}",
]
CXX_FLAGS = ["-O3", "-std=c++17", "-fPIC", "-pthread", "-march=x86-64-v3", f"-I{INCLUDE}", f"-I{CSRC}"]


def _digest(paths) -> str:
    h = hashlib.sha256()
    for p in sorted(paths):
        h.update(str(p.relative_to(ROOT)).encode())
        h.update(p.read_bytes())
    return h.hexdigest()


def _run(cmd):
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("command failed: %s\n%s\n%s" % (" ".join(map(str, cmd)), r.stdout, r.stderr))
    return r.stdout + r.stderr


def _headers():
    return list(CSRC.glob("*.h")) + list(CSRC.glob("*.cuh")) + list(INCLUDE.glob("*.h")) + list((CSRC / "host").glob("*.h"))


def build_cuda(force: bool = False, verbose: bool = False) -> Path:
    out = LIBDIR / "libbigdl_b200.so"
    srcs = sorted(CSRC.glob("*.cu"))
    stamp = LIBDIR / ".cuda.sha256"
    digest = _digest(srcs + _headers()) + "|" + " ".join(NVCC_FLAGS)
    if not force and out.exists() and stamp.exists() and stamp.read_text() == digest:
        return out
    LIBDIR.mkdir(parents=True, exist_ok=True)
    OBJDIR.mkdir(parents=True, exist_ok=True)

    def compile_one(src: Path) -> Path:
        obj = OBJDIR / (src.stem + ".o")
        log = _run([NVCC, *NVCC_FLAGS, "-Xptxas", "-v", "-c", str(src), "-o", str(obj)])
        (OBJDIR / (src.stem + ".ptxas.txt")).write_text(log)
        if verbose:
            print(log)
        return obj

    with ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        objs = list(ex.map(compile_one, srcs))
    _run([NVCC, "-shared", "-o", str(out), *map(str, objs), "-gencode", "arch=compute_100a,code=sm_100a"])
    stamp.write_text(digest)
    return out


def build_host(force: bool = False) -> Path:
    out = LIBDIR / "libbigdl_b200_host.so"
    srcs = sorted((CSRC / "host").glob("*.cpp"))
    if not srcs:
        return out
    stamp = LIBDIR / ".host.sha256"
    digest = _digest(srcs + _headers()) + "|" + " ".join(CXX_FLAGS)
    if not force and out.exists() and stamp.exists() and stamp.read_text() == digest:
        return out
    LIBDIR.mkdir(parents=True, exist_ok=True)
    _run([CXX, *CXX_FLAGS, "-shared", "-Wl,-Bsymbolic", "-o", str(out), *map(str, srcs)])
    stamp.write_text(digest)
    return out


def build_all(force: bool = False, verbose: bool = False):
    return build_cuda(force, verbose), build_host(force)


if __name__ == "__main__":
    libs = build_all(force="--force" in sys.argv, verbose="-v" in sys.argv)
    for lib in libs:
        print(lib, lib.stat().st_size if lib.exists() else "missing")
