from bigdl_b200.dataset.sentence import *  # noqa: F401,F403
