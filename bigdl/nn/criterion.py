from bigdl_b200.nn.criterion import *  # noqa: F401,F403
from bigdl_b200.nn.criterion import Criterion  # noqa: F401

from bigdl_b200.nn.criterion import Criterion as _Criterion  # noqa: E402

if not hasattr(_Criterion, "of"):
    _Criterion.of = classmethod(lambda cls, jcriterion, bigdl_type="float": jcriterion)
