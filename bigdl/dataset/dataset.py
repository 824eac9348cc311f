from bigdl_b200.dataset.base import DataSet  # noqa: F401
