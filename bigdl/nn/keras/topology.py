from bigdl_b200.nn.keras.topology import KerasModel, Model, Sequential  # noqa: F401
