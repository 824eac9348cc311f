from bigdl_b200.dataset.base import *  # noqa: F401,F403
