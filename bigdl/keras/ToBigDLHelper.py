from bigdl_b200.interop.keras_helpers import (to_bigdl_2d_ordering, to_bigdl_2d_padding, to_bigdl_3d_ordering,  # noqa: F401
                                              to_bigdl_3d_padding, to_bigdl_init, to_bigdl_reg)
