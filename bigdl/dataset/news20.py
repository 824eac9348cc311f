from bigdl_b200.dataset.news20 import *  # noqa: F401,F403
