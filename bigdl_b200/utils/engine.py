"""Engine: process-wide device / rank / stream state.

Reference: S/utils/Engine.scala (``Engine.init``, node/core numbers parsed from the Spark master, thread pools
``Engine.default``/``Engine.model``). The B200 mapping is one process per GPU: ``nodeNumber`` = world size,
``coreNumber`` = SM count of the local GPU (CPU: hardware threads); the thread pools become CUDA streams
(compute / comm / copy) plus a small host pool for data loading.
"""
from __future__ import annotations

import logging
import os
import threading
from concurrent.futures import ThreadPoolExecutor

import torch

from . import config

log = logging.getLogger("bigdl_b200")


class _Engine:
    def __init__(self):
        self._inited = False
        self._lock = threading.Lock()
        self.rank = 0
        self.world_size = 1
        self.local_rank = 0
        self.device = torch.device("cpu")
        self._streams = {}
        self._pool = None
        self._singleton_taken = False

    # ---- reference API names
    def init(self, node_number: int | None = None, core_number: int | None = None, on_spark: bool = False,
             device: str | None = None):
        with self._lock:
            self.rank = int(os.environ.get("RANK", 0))
            self.world_size = int(os.environ.get("WORLD_SIZE", 1)) if node_number is None else node_number
            self.local_rank = int(os.environ.get("LOCAL_RANK", 0))
            if device is not None:
                self.device = torch.device(device)
            elif torch.cuda.is_available():
                torch.cuda.set_device(self.local_rank % torch.cuda.device_count())
                self.device = torch.device("cuda", torch.cuda.current_device())
            else:
                self.device = torch.device("cpu")
            self._core_number = core_number
            self._inited = True
        return self

    def ensure(self):
        if not self._inited:
            self.init()
        return self

    def nodeNumber(self) -> int:
        return self.ensure().world_size

    node_number = nodeNumber

    def coreNumber(self) -> int:
        self.ensure()
        if getattr(self, "_core_number", None):
            return self._core_number
        conf = config.get("bigdl.coreNumber")
        if conf:
            return int(conf)
        if self.device.type == "cuda":
            return torch.cuda.get_device_properties(self.device).multi_processor_count
        try:      # physical cores, like the reference's Engine (hyper-threads do not help the BLAS-bound CPU path)
            from . import native

            return native.physical_cores()
        except Exception:
            return os.cpu_count() or 1

    core_number = coreNumber

    def getEngineType(self) -> str:
        return config.get("bigdl.engineType")

    def is_gpu(self) -> bool:
        return self.ensure().device.type == "cuda"

    def compute_dtype(self) -> torch.dtype:
        if self.is_gpu() and config.get("bigdl.computeDtype") == "bf16":
            return torch.bfloat16
        return torch.float32

    # ---- streams (the GPU analogue of Engine.default / Engine.model pools)
    def stream(self, name: str = "compute"):
        self.ensure()
        if self.device.type != "cuda":
            return None
        if name == "compute":
            return torch.cuda.current_stream(self.device)
        if name not in self._streams:
            self._streams[name] = torch.cuda.Stream(self.device)
        return self._streams[name]

    def default_pool(self) -> ThreadPoolExecutor:
        if self._pool is None:
            self._pool = ThreadPoolExecutor(max_workers=max(2, min(16, os.cpu_count() or 2)),
                                            thread_name_prefix="bigdl-default")
        return self._pool

    # ---- singleton check (S/utils/Engine.scala:266)
    def checkSingleton(self) -> bool:
        if not config.get("bigdl.check.singleton"):
            return True
        with self._lock:
            if self._singleton_taken:
                return False
            self._singleton_taken = True
            return True

    def resetSingleton(self):
        self._singleton_taken = False

    def reset(self):
        self._inited = False
        self._streams.clear()


Engine = _Engine()


def init_engine(**kw):
    """Python-API name (P/util/common.py:420)."""
    return Engine.init(**kw)


def get_node_and_core_number():
    return Engine.nodeNumber(), Engine.coreNumber()


class EngineType:
    """The execution back ends ``bigdl.engineType`` selects (S/utils/Engine.scala:37-41). ``MklBlas`` / ``MklDnn`` are
    the reference's two CPU engines — both are accepted and run the host (oracle) path here; ``B200`` is the sm_100a
    engine and the default on a machine with a GPU."""
    MklBlas, MklDnn, B200 = "mklblas", "mkldnn", "b200"


MklBlas, MklDnn = EngineType.MklBlas, EngineType.MklDnn
