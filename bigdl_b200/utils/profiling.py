"""Tracing helpers (the reference has per-module wall-clock ``getTimes()`` and phase timers in ``Metrics``).

* ``enable_nvtx()``: every layer call is wrapped in an NVTX range ``<Class>:<name>`` so a timeline tool shows the model
  structure over the kernels (config key ``bigdl.nvtx`` / env ``BIGDL_NVTX`` turns it on at import).
* ``profile_layers(model, input)``: device time of every leaf layer's forward, measured with CUDA events on the launching
  stream (host wall-clock on CPU) — the per-layer table ``getTimesGroupByModuleType`` gives, but in GPU time.
"""
from __future__ import annotations

import time

import torch

from . import config


def enable_nvtx(on: bool = True) -> None:
    from ..nn import module as M

    M._NVTX[0] = bool(on) and torch.cuda.is_available()


def profile_layers(model, input, warmup: int = 1, iters: int = 3):
    """[(layer name, layer type, mean forward ms)] for every leaf module, in execution order of the first call."""
    leaves = [m for m in model.all_modules() if not m.modules]
    cuda = any(p.is_cuda for p in model.torch_parameters()) if model.torch_parameters() else False
    records: dict[int, list] = {}
    order: list = []

    def wrap(m):
        orig = m.updateOutput

        def timed(x, _m=m, _orig=orig):
            if cuda:
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                y = _orig(x)
                e1.record()
                records.setdefault(id(_m), []).append((e0, e1))
            else:
                t0 = time.perf_counter()
                y = _orig(x)
                records.setdefault(id(_m), []).append(time.perf_counter() - t0)
            if id(_m) not in [id(o) for o in order]:
                order.append(_m)
            return y
        m.__dict__["updateOutput"] = timed

    was_training = model.isTraining()
    model.evaluate()
    try:
        for _ in range(warmup):
            model.forward(input)
        for m in leaves:
            wrap(m)
        for _ in range(iters):
            model.forward(input)
        if cuda:
            torch.cuda.synchronize()
    finally:
        for m in leaves:
            m.__dict__.pop("updateOutput", None)
        if was_training:
            model.training()
    out = []
    for m in order:
        rec = records.get(id(m), [])
        ms = [a.elapsed_time(b) for a, b in rec] if cuda else [r * 1e3 for r in rec]
        out.append((m.getName(), type(m).__name__, sum(ms) / max(len(ms), 1)))
    return out


def format_profile(rows) -> str:
    total = sum(r[2] for r in rows) or 1.0
    lines = [f"{'layer':32s} {'type':28s} {'ms':>9s} {'%':>6s}"]
    for name, typ, ms in rows:
        lines.append(f"{name[:32]:32s} {typ[:28]:28s} {ms:9.4f} {100 * ms / total:6.1f}")
    return "\n".join(lines)


if config.get("bigdl.nvtx", False):
    enable_nvtx(True)
