from bigdl_b200.nn.criterion import *  # noqa: F401,F403
from bigdl_b200.nn.criterion import Criterion  # noqa: F401
