from bigdl_b200.models.vgg import *  # noqa: F401,F403
