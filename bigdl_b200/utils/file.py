"""Object / bytes persistence over local, HDFS and S3 paths (S/utils/File.scala:68-306: ``save``, ``load``,
``saveBytes``, ``readBytes``, ``saveToHdfs``, ``loadFromHdfs``).

Java object serialization becomes pickle (tensors moved to CPU by the callers). Remote schemes are served by Arrow's
native filesystem clients — ``hdfs://`` (libhdfs, needs a Hadoop client install at run time like the reference needs
its jars), ``s3://`` / ``s3a://`` / ``s3n://`` (credentials from ``AWS_ACCESS_KEY_ID`` / ``AWS_SECRET_ACCESS_KEY``, the
variables S/utils/File.scala:112-121 reads into ``fs.s3a.*``), ``gs://``, ``abfs://`` — or by an opener registered with
``register_filesystem`` (tests, custom stores). The overwrite rule is the reference's: an existing target is an error
unless ``isOverwrite``.
"""
from __future__ import annotations

import os
import pickle

_SCHEMES: dict = {}        # scheme -> opener(path, mode) -> file object
_ARROW_FS: dict = {}       # scheme -> (pyarrow FileSystem, strips_scheme)

hdfsPrefix = "hdfs:"
s3aPrefix = "s3a:"


def register_filesystem(scheme: str, opener):
    """``opener(path, mode) -> file object`` for a scheme; or a ``pyarrow.fs.FileSystem`` instance to serve it."""
    try:
        import pyarrow.fs as pfs
        if isinstance(opener, pfs.FileSystem):
            _ARROW_FS[scheme] = opener
            return
    except ImportError:  # pragma: no cover
        pass
    _SCHEMES[scheme] = opener


def _arrow_fs(path: str):
    """(filesystem, path inside it) for a remote URI."""
    import pyarrow.fs as pfs

    scheme, rest = path.split("://", 1)
    if scheme in _ARROW_FS:
        return _ARROW_FS[scheme], rest
    uri = path
    if scheme in ("s3a", "s3n"):          # Hadoop's connector names; Arrow's client speaks plain s3
        uri = "s3://" + rest
    try:
        return pfs.FileSystem.from_uri(uri)
    except Exception as e:
        raise IOError(f"cannot open filesystem for '{path}': {e}") from e


def _exists(path: str) -> bool:
    if "://" not in path:
        return os.path.exists(path)
    scheme = path.split("://", 1)[0]
    if scheme == "file":
        return os.path.exists(path.split("://", 1)[1])
    if scheme in _SCHEMES:
        return False                       # custom openers own their overwrite policy
    import pyarrow.fs as pfs
    fs, p = _arrow_fs(path)
    return fs.get_file_info(p).type != pfs.FileType.NotFound


def _open(path: str, mode: str):
    if "://" in path:
        scheme = path.split("://", 1)[0]
        if scheme == "file":
            path = path.split("://", 1)[1]
        elif scheme in _SCHEMES:
            return _SCHEMES[scheme](path, mode)
        else:
            fs, p = _arrow_fs(path)
            if "w" in mode:
                parent = os.path.dirname(p)
                if parent:
                    try:
                        fs.create_dir(parent, recursive=True)
                    except Exception:
                        pass               # object stores have no directories
                return fs.open_output_stream(p)
            return fs.open_input_stream(p)
    if "w" in mode:
        d = os.path.dirname(os.path.abspath(path))
        os.makedirs(d, exist_ok=True)
    return open(path, mode)


class File:
    hdfsPrefix, s3aPrefix = hdfsPrefix, s3aPrefix

    @staticmethod
    def _check_overwrite(fileName: str, isOverwrite: bool):
        if not isOverwrite and _exists(fileName):
            raise FileExistsError(f"file {fileName} already exists")

    @staticmethod
    def save(obj, fileName: str, isOverwrite: bool = False):
        File._check_overwrite(fileName, isOverwrite)
        with _open(fileName, "wb") as f:
            f.write(pickle.dumps(obj, protocol=pickle.HIGHEST_PROTOCOL))

    @staticmethod
    def load(fileName: str):
        with _open(fileName, "rb") as f:
            return pickle.loads(f.read())

    @staticmethod
    def saveBytes(data: bytes, fileName: str, isOverwrite: bool = False):
        File._check_overwrite(fileName, isOverwrite)
        with _open(fileName, "wb") as f:
            f.write(data)

    @staticmethod
    def readBytes(fileName: str) -> bytes:
        with _open(fileName, "rb") as f:
            return f.read()

    @staticmethod
    def saveToHdfs(obj, fileName: str, overwrite: bool):
        if not fileName.startswith(hdfsPrefix):
            raise ValueError(f"hdfs path {fileName} should have prefix 'hdfs:'")
        File.save(obj, fileName, overwrite)

    @staticmethod
    def loadFromHdfs(fileName: str):
        if not fileName.startswith(hdfsPrefix):
            raise ValueError(f"hdfs path {fileName} should have prefix 'hdfs:'")
        return File.load(fileName)

    @staticmethod
    def exists(fileName: str) -> bool:
        return _exists(fileName)

    save_bytes, read_bytes = saveBytes, readBytes
    register_filesystem = staticmethod(register_filesystem)
