"""``bigdl_b200.nn.layer`` — every layer under one roof, mirroring P/nn/layer.py of the reference."""
from .module import *            # noqa: F401,F403
from .module import Module as Layer   # noqa: F401  (P/nn/layer.py:117 calls the base class ``Layer``)
from .layers_core import *       # noqa: F401,F403
from .layers_table import *      # noqa: F401,F403
from .layers_activation import * # noqa: F401,F403
from .layers_linear import *     # noqa: F401,F403
from .layers_conv_extra import * # noqa: F401,F403
from .layers_recurrent import *  # noqa: F401,F403
from .layers_attention import *  # noqa: F401,F403
from .layers_detection import *  # noqa: F401,F403
from .initialization_method import *  # noqa: F401,F403
from .layers_recurrent import RnnCell as RNN  # noqa: F401  (Python API name, P/nn/layer.py)
from .dynamic_graph import DynamicGraph, FrameManager, Scheduler  # noqa: F401,E402
from .utils import ErrorInfo, NNPrimitive, Utils  # noqa: F401,E402
from .layers_attention import TransformerOperation  # noqa: F401,E402
from . import int8_convertible as _int8_convertible  # noqa: F401,E402  (attaches calcScales & friends to Module)
