from bigdl_b200.dataset.transformer import *  # noqa: F401,F403
