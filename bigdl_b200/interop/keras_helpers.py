"""Small translators from Keras-1.2.2 argument conventions to this framework's (P/keras/ToBigDLHelper.py) and from a
compiled Keras model's loss / optimizer / metrics to criterions, OptimMethods and ValidationMethods
(P/keras/optimization.py ``OptimConverter``). Keras itself is not needed: optimizers and losses are recognised by
name, by config dict or by duck-typed attributes (``lr``, ``decay``, ``momentum``, ...)."""
from __future__ import annotations

import math
import warnings


def to_bigdl_2d_ordering(order: str) -> str:
    try:
        return {"tf": "NHWC", "th": "NCHW"}[order]
    except KeyError:
        raise Exception(f"Unsupported dim_ordering: {order}") from None


def to_bigdl_3d_ordering(order: str) -> str:
    try:
        return {"tf": "channel_last", "th": "channel_first"}[order]
    except KeyError:
        raise Exception(f"Unsupported dim_ordering: {order}") from None


def to_bigdl_3d_padding(border_mode: str):
    if border_mode == "valid":
        return 0, 0, 0
    if border_mode == "same":
        return -1, -1, -1
    raise Exception(f"Unsupported border mode: {border_mode}")


def _same_pad(x, k, d, dilation):
    return int(math.ceil((x * (d - 1) + dilation * (k - 1) - d + 1) / 2))


def to_bigdl_2d_padding(border_mode: str, *args):
    """``same`` -> (-1, -1) (the layers compute it), or explicit pads from (h, kh, dh, dilation_h[, w, kw, dw, dilation_w])."""
    if border_mode == "valid":
        return 0, 0
    if border_mode != "same":
        raise Exception(f"Unsupported border mode: {border_mode}")
    if not args:
        return -1, -1
    if len(args) == 4:
        return _same_pad(*args), 0
    if len(args) == 8:
        return _same_pad(*args[:4]), _same_pad(*args[4:])
    raise Exception("to_bigdl_2d_padding takes 0, 4 or 8 geometry arguments")


def to_bigdl_init(kinit_method: str):
    from ..nn import initialization_method as I

    table = {"glorot_uniform": lambda: I.Xavier(), "one": lambda: I.Ones(), "zero": lambda: I.Zeros(),
             "uniform": lambda: I.RandomUniform(-0.05, 0.05), "normal": lambda: I.RandomNormal(0.0, 0.05)}
    if kinit_method not in table:
        raise Exception(f"Unsupported init type: {kinit_method}")
    return table[kinit_method]()


def to_bigdl_reg(reg):
    """Keras regularizer config ``{"l1": a, "l2": b}`` -> ``L1L2Regularizer`` (None stays None)."""
    if not reg:
        return None
    from ..optim.regularizer import L1L2Regularizer

    return L1L2Regularizer(float(reg.get("l1", 0.0)), float(reg.get("l2", 0.0)))


def _scalar(v, default=0.0):
    if v is None:
        return default
    for attr in ("numpy", "item", "get_value", "eval"):
        f = getattr(v, attr, None)
        if callable(f):
            try:
                return float(f())
            except Exception:
                continue
    return float(v)


class OptimConverter:
    @staticmethod
    def to_bigdl_metrics(metrics):
        from ..optim import validation as V

        out = []
        for m in (metrics if isinstance(metrics, (list, tuple)) else [metrics]):
            name = m if isinstance(m, str) else getattr(m, "__name__", str(m))
            if name in ("accuracy", "acc"):
                out.append(V.Top1Accuracy())
            elif name in ("top_k_categorical_accuracy", "top5"):
                out.append(V.Top5Accuracy())
            elif name in ("mae", "mean_absolute_error"):
                out.append(V.MAE())
            elif name == "loss":
                out.append(V.Loss())
            else:
                raise Exception(f"Unsupported metric: {name}")
        return out

    @staticmethod
    def to_bigdl_criterion(kloss):
        from ..nn import criterion as C

        name = kloss.lower() if isinstance(kloss, str) else getattr(kloss, "__name__", str(kloss)).lower()
        table = {
            "categorical_crossentropy": lambda: C.CategoricalCrossEntropy(),
            "mse": lambda: C.MSECriterion(), "mean_squared_error": lambda: C.MSECriterion(),
            "binary_crossentropy": lambda: C.BCECriterion(),
            "mae": lambda: C.AbsCriterion(), "mean_absolute_error": lambda: C.AbsCriterion(),
            "hinge": lambda: C.MarginCriterion(),
            "squared_hinge": lambda: C.MarginCriterion(squared=True),
            "mape": lambda: C.MeanAbsolutePercentageCriterion(),
            "mean_absolute_percentage_error": lambda: C.MeanAbsolutePercentageCriterion(),
            "msle": lambda: C.MeanSquaredLogarithmicCriterion(),
            "mean_squared_logarithmic_error": lambda: C.MeanSquaredLogarithmicCriterion(),
            "sparse_categorical_crossentropy": lambda: C.ClassNLLCriterion(logProbAsInput=False),
            "kld": lambda: C.KullbackLeiblerDivergenceCriterion(),
            "kullback_leibler_divergence": lambda: C.KullbackLeiblerDivergenceCriterion(),
            "poisson": lambda: C.PoissonCriterion(),
            "cosine": lambda: C.CosineProximityCriterion(), "cosine_proximity": lambda: C.CosineProximityCriterion(),
        }
        if name not in table:
            raise Exception(f"Not supported loss: {kloss}")
        return table[name]()

    @staticmethod
    def to_bigdl_optim_method(koptim):
        """``koptim``: a Keras optimizer object, its ``get_config()`` dict (with a ``name`` / ``class_name`` entry) or a
        name such as "sgd"."""
        from ..optim import optim_method as O

        if isinstance(koptim, str):
            kind, get = koptim.lower(), (lambda k, d=None: d)
        elif isinstance(koptim, dict):
            cfg = dict(koptim.get("config", koptim))
            kind = str(koptim.get("class_name", koptim.get("name", ""))).lower()
            get = lambda k, d=None: cfg.get(k, d)
        else:
            kind = type(koptim).__name__.lower()
            get = lambda k, d=None: getattr(koptim, k, d)
        lr, decay = _scalar(get("lr"), 0.01), _scalar(get("decay"), 0.0)
        if kind == "sgd":
            return O.SGD(learningrate=lr, learningrate_decay=decay, momentum=_scalar(get("momentum"), 0.0),
                         dampening=0.0, nesterov=bool(get("nesterov", False)))
        if kind == "adagrad":
            warnings.warn("For Adagrad, we don't support epsilon for now")
            return O.Adagrad(learningrate=lr, learningrate_decay=decay)
        if kind == "adam":
            return O.Adam(learningrate=_scalar(get("lr"), 0.001), learningrate_decay=decay,
                          beta1=_scalar(get("beta_1"), 0.9), beta2=_scalar(get("beta_2"), 0.999),
                          epsilon=_scalar(get("epsilon"), 1e-8))
        if kind == "rmsprop":
            return O.RMSprop(learningrate=_scalar(get("lr"), 0.001), learningrate_decay=decay,
                             decayrate=_scalar(get("rho"), 0.9), epsilon=_scalar(get("epsilon"), 1e-8))
        if kind == "adadelta":
            warnings.warn("For Adadelta, we don't support learning rate and learning rate decay for now")
            return O.Adadelta(decayrate=_scalar(get("rho"), 0.95), epsilon=_scalar(get("epsilon"), 1e-8))
        if kind == "adamax":
            warnings.warn("For Adamax, we don't support learning rate decay for now")
            return O.Adamax(learningrate=_scalar(get("lr"), 0.002), beta1=_scalar(get("beta_1"), 0.9),
                            beta2=_scalar(get("beta_2"), 0.999), epsilon=_scalar(get("epsilon"), 1e-8))
        raise Exception(f"Unsupported optimizer: {koptim}")
