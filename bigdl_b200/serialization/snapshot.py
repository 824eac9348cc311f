"""Object snapshots: ``model.save(path)`` / ``Module.load(path)`` and ``OptimMethod.save/load``.

The reference writes these with *Java object serialization* (S/utils/File.scala:68,176; used for checkpoints by
S/optim/Optimizer.scala:548-586), which cannot be produced or read without a JVM. The equivalent here is a Python
object snapshot (``torch.save``) with the same file naming (``model.<neval>``, ``optimMethod-<name>.<neval>``,
``state.<neval>``) and the same content rules: ``OptimMethod.save`` drops the history buffers and keeps the state table
(S/optim/OptimMethod.scala:92-96). The portable format is the protobuf ``.bigdl`` file (``saveModule``/``loadModule``).
"""
from __future__ import annotations

import copy
import glob
import os

import torch


def _check_overwrite(path: str, overwrite: bool):
    if os.path.exists(path) and not overwrite:
        raise FileExistsError(f"{path} already exists (pass overWrite=True / call overWriteCheckpoint())")
    d = os.path.dirname(os.path.abspath(path))
    os.makedirs(d, exist_ok=True)


def _portable_copy(module):
    c = module.cloneModule()
    c.to("cpu")
    c.clearState()
    for m in c.all_modules():
        m.__dict__.pop("_param_sync_keepalive", None)
        m.__dict__.pop("_param_sync", None)
    return c


def save_snapshot(module, path: str, overwrite: bool = False):
    _check_overwrite(path, overwrite)
    tmp = path + ".tmp"
    torch.save({"format": "bigdl_b200.module.v1", "module": _portable_copy(module)}, tmp)
    os.replace(tmp, path)


def load_snapshot(path: str):
    obj = torch.load(path, map_location="cpu", weights_only=False)
    if not isinstance(obj, dict) or obj.get("format") != "bigdl_b200.module.v1":
        raise ValueError(f"{path} is not a bigdl_b200 module snapshot")
    return obj["module"]


def load_snapshot_into(module, path: str):
    """Copy the weights (and buffers) of a snapshot into an existing, structurally identical model."""
    src = load_snapshot(path)
    sp, dp = src.torch_parameters(), module.torch_parameters()
    if len(sp) != len(dp):
        raise ValueError("snapshot and model have different parameter lists")
    with torch.no_grad():
        for s, d in zip(sp, dp):
            d.copy_(s.to(d.device))
        for sm, dm in zip(src.all_modules(), module.all_modules()):
            for k, b in sm._buffers.items():
                if k in dm._buffers and isinstance(b, torch.Tensor):
                    dm._buffers[k].copy_(b.to(dm._buffers[k].device))
    module._mark_weights_dirty()
    return module


def save_optim_method(om, path: str, overwrite: bool = False):
    _check_overwrite(path, overwrite)
    c = copy.copy(om)
    c._hist = {}                      # clearHistory(): momentum / moment buffers are not persisted
    c.state = om.state.clone()
    for k in ("old_dirs", "old_stps", "g_old", "dir"):
        c.state.delete(k)
    tmp = path + ".tmp"
    try:
        torch.save({"format": "bigdl_b200.optim.v1", "optim": c}, tmp)
    except (AttributeError, TypeError) as e:   # pickle's complaint about a lambda / local function
        if os.path.exists(tmp):
            os.remove(tmp)
        raise TypeError(f"optim method {type(om).__name__} cannot be saved: {e}. Learning-rate schedules must hold "
                        "module-level functions or callable objects (e.g. optim_method.EpochBoundaries), not "
                        "lambdas or local functions") from e
    os.replace(tmp, path)


def load_optim_method(path: str):
    obj = torch.load(path, map_location="cpu", weights_only=False)
    if not isinstance(obj, dict) or obj.get("format") != "bigdl_b200.optim.v1":
        raise ValueError(f"{path} is not a bigdl_b200 optim-method snapshot")
    return obj["optim"]


def latest_file(root: str, prefix: str) -> str | None:
    """Newest ``<prefix>*`` file by modification time, searching the timestamped sub-directories too
    (S/optim/DistriOptimizer.scala:986-1003 getLatestFile)."""
    cands = glob.glob(os.path.join(root, prefix + "*")) + glob.glob(os.path.join(root, "*", prefix + "*"))
    cands = [c for c in cands if os.path.isfile(c) and not c.endswith(".tmp")]
    if not cands:
        return None
    return max(cands, key=lambda p: (os.path.getmtime(p), p))
