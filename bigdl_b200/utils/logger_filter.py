"""Log routing (S/utils/LoggerFilter.scala:90-134): framework INFO lines stay on the console, chatty third-party loggers
(Spark/Akka/Hadoop in the reference; torch.distributed / NCCL bootstrap / filelock here) are redirected to ``bigdl.log``.
Switches: ``bigdl.utils.LoggerFilter.disable``, ``.logFile``, ``.enableSparkLog``."""
from __future__ import annotations

import logging
import os

from . import config

_NOISY = ("torch.distributed", "torch.nn.parallel", "filelock", "urllib3", "py4j", "org.apache.spark")
_installed = False


def redirectSparkInfoLogs(log_file: str | None = None):
    global _installed
    if config.get("bigdl.utils.LoggerFilter.disable") or _installed:
        return
    path = log_file or config.get("bigdl.utils.LoggerFilter.logFile") or "bigdl.log"
    if not os.path.isabs(path):
        path = os.path.join(os.getcwd(), path)
    fmt = logging.Formatter("%(asctime)s %(levelname)-5s %(name)s:%(lineno)d - %(message)s")
    console = logging.StreamHandler()
    console.setFormatter(fmt)
    console.setLevel(logging.INFO)
    fw = logging.getLogger("bigdl_b200")
    fw.setLevel(logging.INFO)
    if not fw.handlers:
        fw.addHandler(console)
    fw.propagate = False
    try:
        fh = logging.FileHandler(path)
        fh.setFormatter(fmt)
        for name in _NOISY:
            lg = logging.getLogger(name)
            lg.addHandler(fh)
            lg.propagate = False
            lg.setLevel(logging.INFO if config.get("bigdl.utils.LoggerFilter.enableSparkLog") else logging.ERROR)
    except OSError:
        pass
    _installed = True


redire_spark_logs = redirectSparkInfoLogs


def show_bigdl_info_logs(level=logging.INFO):
    logging.getLogger("bigdl_b200").setLevel(level)


class LoggerFilter:
    """Object form of the two functions above (S/utils/LoggerFilter.scala:90-134)."""
    redirectSparkInfoLogs = staticmethod(redirectSparkInfoLogs)
    show_bigdl_info_logs = staticmethod(show_bigdl_info_logs)
