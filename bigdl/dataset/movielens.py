from bigdl_b200.dataset.movielens import *  # noqa: F401,F403
