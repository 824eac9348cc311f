"""Object / bytes persistence (S/utils/File.scala:68-306: ``save``, ``load``, ``saveBytes``, ``readBytes`` over local,
HDFS and S3 paths). Java object serialization becomes pickle with tensors moved to CPU; remote schemes are resolved
through ``fsspec``-style hooks when one is registered, else rejected with a clear error (no network in this build)."""
from __future__ import annotations

import os
import pickle

_SCHEMES: dict = {}


def register_filesystem(scheme: str, opener):
    """``opener(path, mode) -> file object`` for e.g. "hdfs" or "s3a"."""
    _SCHEMES[scheme] = opener


def _open(path: str, mode: str):
    if "://" in path:
        scheme = path.split("://", 1)[0]
        if scheme == "file":
            path = path.split("://", 1)[1]
        elif scheme in _SCHEMES:
            return _SCHEMES[scheme](path, mode)
        else:
            raise IOError(f"no filesystem registered for scheme '{scheme}' (File.register_filesystem)")
    if "w" in mode:
        d = os.path.dirname(os.path.abspath(path))
        os.makedirs(d, exist_ok=True)
    return open(path, mode)


class File:
    @staticmethod
    def save(obj, fileName: str, isOverwrite: bool = False):
        if not isOverwrite and "://" not in fileName and os.path.exists(fileName):
            raise FileExistsError(f"file {fileName} already exists")
        with _open(fileName, "wb") as f:
            pickle.dump(obj, f, protocol=pickle.HIGHEST_PROTOCOL)

    @staticmethod
    def load(fileName: str):
        with _open(fileName, "rb") as f:
            return pickle.load(f)

    @staticmethod
    def saveBytes(data: bytes, fileName: str, isOverwrite: bool = False):
        if not isOverwrite and "://" not in fileName and os.path.exists(fileName):
            raise FileExistsError(f"file {fileName} already exists")
        with _open(fileName, "wb") as f:
            f.write(data)

    @staticmethod
    def readBytes(fileName: str) -> bytes:
        with _open(fileName, "rb") as f:
            return f.read()

    save_bytes, read_bytes = saveBytes, readBytes
    register_filesystem = staticmethod(register_filesystem)
