from bigdl_b200.models.autoencoder import *  # noqa: F401,F403
