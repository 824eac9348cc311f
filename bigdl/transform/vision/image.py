from bigdl_b200.transform.vision.image import *  # noqa: F401,F403
