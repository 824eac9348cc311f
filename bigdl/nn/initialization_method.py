from bigdl_b200.nn.initialization_method import *  # noqa: F401,F403
