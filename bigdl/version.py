from bigdl_b200 import __version__  # noqa: F401
