from bigdl_b200.dlframes.dl_image import DLImageReader  # noqa: F401
