from bigdl_b200.models.lenet import LeNet5 as build_model  # noqa: F401
