from bigdl_b200.dlframes.dl_image import DLImageTransformer  # noqa: F401
