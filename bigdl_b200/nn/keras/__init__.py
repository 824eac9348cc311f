"""Keras-1.2.2-style API (S/nn/keras, P/nn/keras)."""
from .layers import *  # noqa: F401,F403
from .layers import Input, KerasLayer, merge
from .topology import KerasModel, Model, Sequential

layer = __import__("bigdl_b200.nn.keras.layers", fromlist=["*"])
topology = __import__("bigdl_b200.nn.keras.topology", fromlist=["*"])
