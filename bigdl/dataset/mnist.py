from bigdl_b200.dataset.mnist import *  # noqa: F401,F403
