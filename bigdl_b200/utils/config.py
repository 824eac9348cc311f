"""Typed configuration registry.

The reference reads JVM system properties ad hoc (``-Dbigdl.engineType=...``; full list in SURVEY.md §5.6). Here one
registry owns every flag: values come from (highest priority first) ``set()``, environment variables
(``BIGDL_<NAME>`` with dots replaced by underscores, upper-cased), then the defaults below. The old property names
are accepted verbatim so launch scripts written for the reference keep working.
"""
from __future__ import annotations

import os
from typing import Any

_DEFAULTS: dict[str, Any] = {
    # engine
    "bigdl.engineType": "b200",           # reference: mklblas | mkldnn (S/utils/Engine.scala:210-216)
    "bigdl.localMode": True,
    "bigdl.coreNumber": 0,                # 0 = autodetect
    "bigdl.check.singleton": False,
    "bigdl.computeDtype": "bf16",         # bf16 on GPU, fp32 on CPU
    "bigdl.enableNHWC": True,
    # failure / retry (S/optim/DistriOptimizer.scala:881-963)
    "bigdl.failure.retryTimes": 5,
    "bigdl.failure.retryTimeInterval": 120,
    # parameter sync
    "bigdl.Parameter.syncPoolSize": 4,
    "bigdl.parameter.compress": "bf16_rn",   # bf16_rn | bf16_trunc (reference-exact) | fp32
    "bigdl.parameter.backend": "auto",       # auto | p2p (fused NVLink kernels) | nccl | gloo
    "bigdl.parameter.p2p": "bucketed",       # bucketed (exchange overlapped with backward, pulled all-gather) | monolithic
    "bigdl.parameter.overlap": True,         # False: one bucket = the whole vector (no overlap; for A/B measurements)
    "bigdl.parameter.bucketMB": 8.0,         # fp32 gradient bytes per bucket of the bucketed p2p backend
    "bigdl.parameter.firstBucketMB": 1.0,    # bucket 0 cannot overlap backward (its input has no gradient): keep it small
    "bigdl.parameter.commBlocks": 96,        # CTAs per communication kernel launched INSIDE backward (they take SMs from it)
    "bigdl.parameter.allgather": "push",     # push (owner stores to every peer after its update) | pull (consumers fetch)
    # Linear weights at least this big (bf16 MB) are NOT pushed by their owners; the forward GEMM gathers the remote row
    # blocks itself (plain TMA loads from the owners' memory, written through to the local copy). 0 = off: measured on
    # VGG-16 the pushed all-gather hides behind backward, the fused one sits on the forward's critical path
    # (profiles/comm_r2.md), so it is opt-in.
    "bigdl.parameter.gatherGemmMB": 0.0,
    "bigdl.parameter.nvls": False,           # ParallelOptimizer buckets: all-reduce inside the NVSwitch (multimem.ld_reduce, csrc/nvls.cu)
    "bigdl.parallelOptimizer.iterationPerTime": 1,
    "bigdl.parallelOptimizer.parameterBlocks": 10,
    "bigdl.parallel.bucketMB": 16.0,         # ParallelOptimizer bucket size (fp32 gradient bytes)
    # fusion flags (S/nn/mkldnn/DnnGraph.scala:412-463)
    "bigdl.mkldnn.fusion": True,
    "bigdl.mkldnn.fusion.convbn": True,
    "bigdl.mkldnn.fusion.bnrelu": True,
    "bigdl.mkldnn.fusion.convrelu": True,
    "bigdl.mkldnn.fusion.convsum": True,
    "bigdl.fusion.bnadd": True,              # BatchNorm closing a residual branch + shortcut add + ReLU as one kernel (fwd and bwd)
    "bigdl.fusion.bnstats": True,            # conv followed by BatchNorm: sum / sum of squares from the conv's GEMM epilogue
    # conv+ReLU -> conv: the first conv's ReLU backward and bias-gradient sums run in the second conv's dgrad epilogue.
    # Off by default: measured slower than the separate mask + column-sum pass (VGG-16 18.43 vs 18.29 ms, Inception-v1
    # 5.49 vs 5.43 ms): those dgrad GEMMs are epilogue-bound, the pass it saves runs at HBM speed.
    "bigdl.fusion.reludgrad": False,
    "bigdl.fusion.skipgrad": True,           # residual block input: sum its two gradients in a dgrad GEMM epilogue
    "bigdl.fusion.residual": True,           # inference: shortcut add + ReLU inside the last conv's GEMM epilogue
    # misc
    "bigdl.quantized.blockScale": False,     # Quantizer: MX block-scaled fp8 (one scale per 32 inputs) for Linear layers
    "bigdl.metrics": False,                  # per-phase host timers in the training loop (Metrics summary at the end)
    "bigdl.rnn.persistent": True,            # LSTM layers: one persistent kernel per sequence (csrc/lstm_seq.cu)
    "bigdl.cudaGraph": True,
    "bigdl.nvtx": False,                     # NVTX range per layer call (utils/profiling.py)
    "bigdl.wgrad.stream": False,             # issue weight-gradient GEMMs on a side stream (joined after backward)
    "bigdl.concat.streams": True,            # run the branches of a channel Concat on side streams
    "bigdl.utils.LoggerFilter.disable": False,
    "bigdl.utils.LoggerFilter.logFile": "bigdl.log",
    "bigdl.utils.LoggerFilter.enableSparkLog": True,
    "bigdl.tensor.fold": 1000,
    "bigdl.tensor.nativeMath": "cuda",       # fp32/fp64 Tensor BLAS + VML on the native entry points: cuda | all | off
    "bigdl.ModelBroadcastFactory": "default",
}

_overrides: dict[str, Any] = {}


def _env_name(key: str) -> str:
    return key.replace(".", "_").upper()


def _coerce(val: str, like: Any):
    if isinstance(like, bool):
        return val.strip().lower() in ("1", "true", "yes", "on")
    if isinstance(like, int):
        return int(val)
    if isinstance(like, float):
        return float(val)
    return val


def get(key: str, default: Any = None) -> Any:
    if key in _overrides:
        return _overrides[key]
    base = _DEFAULTS.get(key, default)
    env = os.environ.get(_env_name(key))
    if env is None:
        env = os.environ.get(key)  # exact property name also accepted
    if env is not None:
        return _coerce(env, base) if base is not None else env
    return base


def set(key: str, value: Any) -> None:  # noqa: A001 - mirrors System.setProperty
    _overrides[key] = value


def unset(key: str) -> None:
    _overrides.pop(key, None)


def keys():
    return sorted(set_union(_DEFAULTS.keys(), _overrides.keys()))


def set_union(a, b):
    out = dict.fromkeys(a)
    out.update(dict.fromkeys(b))
    return out.keys()


def parse_java_opts(opts: str) -> None:
    """Accept a ``-Dbigdl.x=y -Dbigdl.z=w`` string as used with spark-submit in the reference."""
    for tok in opts.split():
        if tok.startswith("-D") and "=" in tok:
            k, v = tok[2:].split("=", 1)
            like = _DEFAULTS.get(k)
            _overrides[k] = _coerce(v, like) if like is not None else v
