from bigdl_b200.nn.keras.layers import *  # noqa: F401,F403
from bigdl_b200.nn.keras.layers import Input, KerasLayer, merge  # noqa: F401
