"""OptimMethods and learning-rate schedules.

Reference: S/optim/OptimMethod.scala, SGD.scala (+ the 13 LearningRateSchedule classes inside the SGD object,
:233-690), Adam.scala, ParallelAdam.scala, Adagrad.scala, Adadelta.scala, Adamax.scala, RMSprop.scala, Ftrl.scala,
LBFGS.scala (+LineSearch.scala), LarsSGD.scala. Formulas are copied *as math* from those files;
the execution is one fused flat kernel on GPU (csrc/optim.cu) and plain torch ops on CPU.

Quirks preserved: schedules store a NEGATIVE ``currentRate``; SGD's ``dampening`` defaults to ``momentum``;
``OptimMethod.save`` drops the history buffers (``clearHistory``) and keeps only the state table.
"""
from __future__ import annotations

import copy
import ctypes
import math
import struct

import torch

from ..utils.table import Table, T


# ====================================================================== schedules
class LearningRateSchedule:
    def __init__(self):
        self.currentRate = 0.0          # NEGATIVE learning rate, as in the reference
        self.excludeIterations = 0
        self.excludeEpochs = 0
        self.maxIterations = 0

    def updateHyperParameter(self, optim: "SGD"):  # pragma: no cover
        raise NotImplementedError

    update_hyper_parameter = updateHyperParameter


class Default(LearningRateSchedule):
    def updateHyperParameter(self, o):
        n = o.state.getOrElse("evalCounter", 0)
        self.currentRate = -o.learningRate / (1 + (n - self.excludeIterations) * o.learningRateDecay)
        o.state["evalCounter"] = n + 1


class Step(LearningRateSchedule):
    def __init__(self, step_size: int, gamma: float):
        super().__init__()
        self.step_size, self.gamma = step_size, gamma

    def updateHyperParameter(self, o):
        n = o.state.getOrElse("evalCounter", 0)
        k = max((n - self.excludeIterations) // self.step_size, 0)
        self.currentRate = -o.learningRate * (self.gamma ** k)
        o.state["evalCounter"] = n + 1


class MultiStep(LearningRateSchedule):
    def __init__(self, step_sizes, gamma: float):
        super().__init__()
        self.step_sizes, self.gamma = list(step_sizes), gamma

    def updateHyperParameter(self, o):
        n = o.state.getOrElse("evalCounter", 0)
        clr = -o.learningRate
        for s in self.step_sizes:
            if (n - self.excludeIterations) >= s:
                clr *= self.gamma
            else:
                break
        self.currentRate = clr
        o.state["evalCounter"] = n + 1


class EpochBoundaries:
    """Picklable ``epoch -> exponent`` function for EpochDecay / EpochDecayWithWarmUp: the number of boundaries that
    ``epoch`` has reached (``EpochBoundaries(30, 60, 80)`` is TrainImageNet.scala:117-120's imageNetDecay). A local
    function works too, but an OptimMethod holding one cannot be written by setCheckpoint."""

    def __init__(self, *boundaries: int):
        self.boundaries = tuple(sorted(int(b) for b in boundaries))

    def __call__(self, epoch: int) -> int:
        return sum(1 for b in self.boundaries if epoch >= b)


class EpochDecay(LearningRateSchedule):
    def __init__(self, decay_type):
        super().__init__()
        self.decay_type = decay_type

    def updateHyperParameter(self, o):
        epoch = o.state["epoch"]
        self.currentRate = -o.learningRate * math.pow(0.1, self.decay_type(epoch - self.excludeEpochs))


class EpochStep(LearningRateSchedule):
    def __init__(self, step_size: int, gamma: float):
        super().__init__()
        self.step_size, self.gamma = step_size, gamma

    def updateHyperParameter(self, o):
        epoch = o.state["epoch"]
        k = max((epoch - self.excludeEpochs) // self.step_size, 0)
        self.currentRate = -o.learningRate * (self.gamma ** k)


class Regime:
    def __init__(self, start_epoch: int, end_epoch: int, config: Table | dict):
        self.startEpoch, self.endEpoch = start_epoch, end_epoch
        self.config = config if isinstance(config, Table) else Table(**config)


class EpochSchedule(LearningRateSchedule):
    _FIELDS = {"learningRate", "learningRateDecay", "weightDecay", "momentum", "dampening", "nesterov",
               "learningRates", "weightDecays"}

    def __init__(self, regimes):
        super().__init__()
        self.regimes = list(regimes)

    def updateHyperParameter(self, o):
        epoch = o.state["epoch"] - self.excludeEpochs
        for r in self.regimes:
            if r.startEpoch <= epoch <= r.endEpoch:
                for k, v in r.config.items():
                    if k == "leaningRateSchedule":
                        o.learningRateSchedule = v
                    elif k in self._FIELDS:
                        setattr(o, k, v)
                    else:
                        raise ValueError(f"EpochSchedule: {k} is not a member of SGD")
        self.currentRate = -o.learningRate


class Poly(LearningRateSchedule):
    def __init__(self, power: float, max_iteration: int):
        super().__init__()
        self.power, self.max_iteration = power, max_iteration

    def updateHyperParameter(self, o):
        n = o.state.getOrElse("evalCounter", 0)
        self.currentRate = 0.0 if n > self.max_iteration else \
            -o.learningRate * math.pow(1.0 - n / self.max_iteration, self.power)
        o.state["evalCounter"] = n + 1


class NaturalExp(LearningRateSchedule):
    def __init__(self, decay_step: int, gamma: float):
        super().__init__()
        self.decay_step, self.gamma = decay_step, gamma

    def updateHyperParameter(self, o):
        n = o.state.getOrElse("evalCounter", 0)
        p = (n - self.excludeIterations) // self.decay_step
        self.currentRate = -o.learningRate * math.exp(-self.gamma * p)
        o.state["evalCounter"] = n + 1


class Exponential(LearningRateSchedule):
    def __init__(self, decay_step: int, decay_rate: float, stair_case: bool = False):
        super().__init__()
        self.decay_step, self.decay_rate, self.stair_case = decay_step, decay_rate, stair_case

    def updateHyperParameter(self, o):
        n = o.state.getOrElse("evalCounter", 0)
        p = (n - self.excludeIterations) / float(self.decay_step)
        if self.stair_case:
            p = math.floor(p)
        self.currentRate = -o.learningRate * math.pow(self.decay_rate, p)
        o.state["evalCounter"] = n + 1


class Plateau(LearningRateSchedule):
    def __init__(self, monitor: str, factor: float = 0.1, patience: int = 10, mode: str = "min",
                 epsilon: float = 1e-4, cooldown: int = 0, min_lr: float = 0.0):
        super().__init__()
        if factor >= 1:
            raise ValueError("Plateau does not support a factor >= 1.0")
        if mode not in ("min", "max"):
            raise ValueError(f"Learning Rate Plateau Reducing mode {mode} is unknown, please use min | max")
        self.monitor, self.factor, self.patience, self.mode = monitor, factor, patience, mode
        self.epsilon, self.cooldown, self.min_lr = epsilon, cooldown, min_lr
        self.best = float("inf") if mode == "min" else float("-inf")
        self.cooldownCounter = self.waitCounter = 0
        self.curEpoch = 1

    def _better(self, a, b):
        return a < b - self.epsilon if self.mode == "min" else a > b + self.epsilon

    def updateHyperParameter(self, o):
        epoch = o.state["epoch"] - self.excludeEpochs
        if epoch == 1:
            self.currentRate = -o.learningRate
        if epoch == self.curEpoch:
            return
        self.curEpoch = epoch
        cur = o.state.get(self.monitor)
        if cur is None:
            raise ValueError(f"Learning Rate Plateau Reducing requires {self.monitor} available!")
        if self.cooldownCounter > 0:
            self.cooldownCounter -= 1
            self.waitCounter = 0
        if self._better(cur, self.best):
            self.best = cur
            self.waitCounter = 0
        elif self.cooldownCounter <= 0:
            if self.waitCounter >= self.patience:
                if abs(self.currentRate) > self.min_lr + self.min_lr * 1e-4:
                    self.currentRate = -max(abs(self.currentRate) * self.factor, self.min_lr)
                    self.cooldownCounter = self.cooldown
                    self.waitCounter = 0
            self.waitCounter += 1


class Warmup(LearningRateSchedule):
    def __init__(self, delta: float):
        super().__init__()
        self.delta = delta

    def updateHyperParameter(self, o):
        n = o.state.getOrElse("evalCounter", 0)
        self.currentRate = -o.learningRate - self.delta * (n - self.excludeIterations)
        o.state["evalCounter"] = n + 1


class SequentialSchedule(LearningRateSchedule):
    def __init__(self, iteration_per_epoch: int):
        super().__init__()
        self.iteration_per_epoch = iteration_per_epoch
        self.schedules: list[LearningRateSchedule] = []
        self.cur = 0

    def add(self, schedule: LearningRateSchedule, max_iteration: int):
        schedule.excludeIterations = 0 if not self.schedules else self.schedules[-1].maxIterations
        schedule.maxIterations = schedule.excludeIterations + max_iteration
        schedule.excludeEpochs = schedule.excludeIterations // self.iteration_per_epoch
        self.schedules.append(schedule)
        return self

    def updateHyperParameter(self, o):
        n = o.state.getOrElse("evalCounter", 0)
        if n > self.schedules[self.cur].maxIterations and self.cur + 1 < len(self.schedules):
            o.learningRate = -self.currentRate
            self.cur += 1
        self.schedules[self.cur].updateHyperParameter(o)
        self.currentRate = self.schedules[self.cur].currentRate


class EpochDecayWithWarmUp(LearningRateSchedule):
    def __init__(self, warm_up_iteration: int, warm_up_delta: float, decay_type):
        super().__init__()
        self.warm_up_iteration, self.warm_up_delta, self.decay_type = warm_up_iteration, warm_up_delta, decay_type

    def updateHyperParameter(self, o):
        n = o.state.getOrElse("evalCounter", 0)
        if n < self.warm_up_iteration:
            clr = -o.learningRate - self.warm_up_delta * n
        else:
            max_lr = o.learningRate + self.warm_up_delta * self.warm_up_iteration
            clr = -max_lr * math.pow(0.1, self.decay_type(o.state["epoch"]))
        o.state["evalCounter"] = n + 1
        self.currentRate = clr


# ====================================================================== native flat-update plumbing
KIND_SGD, KIND_ADAM, KIND_ADAGRAD, KIND_RMSPROP, KIND_ADADELTA, KIND_ADAMAX, KIND_FTRL, KIND_LARS = range(8)


def pack_hp(lr=0.0, wd=0.0, momentum=0.0, dampening=0.0, eps=0.0, gscale=1.0, clip_min=0.0, clip_max=0.0,
            nesterov=0, first_step=0, lr_power=0.0, l1=0.0, l2=0.0, l2_shrink=0.0) -> bytes:
    """Byte image of ``struct OptimHP`` (csrc/optim_update.cuh)."""
    return struct.pack("<8f2i4f", lr, wd, momentum, dampening, eps, gscale, clip_min, clip_max, int(nesterov),
                       int(first_step), lr_power, l1, l2, l2_shrink)


def native_update(kind: int, w: torch.Tensor, g: torch.Tensor, s0, s1, hp: bytes | None, w16=None, truncate=False,
                  zero_grad=False, hp_dev: int | None = None):
    """Flat fused update. ``hp`` (host bytes) is passed by value; ``hp_dev`` (device address of an OptimHP) wins when
    given — that is what a captured CUDA graph uses so the schedule can change between replays."""
    from ..ops import _lib

    lib = _lib.cuda_lib()
    buf = ctypes.create_string_buffer(hp, len(hp)) if hp is not None else None
    _lib.check(lib.bdl_optim_update(kind, w.data_ptr(), g.data_ptr(), _lib.ptr(s0), _lib.ptr(s1), _lib.ptr(w16),
                                    w.numel(), ctypes.cast(buf, ctypes.c_void_p) if buf is not None else None, hp_dev,
                                    int(truncate), int(zero_grad), _lib.stream_ptr()), "bdl_optim_update")
    _lib.count_launch()


def run_plan(plan, x, g, w16=None, truncate=False, zero_grad=False):
    kind, hp, s0, s1 = plan
    native_update(kind, x, g, s0, s1, hp, w16, truncate, zero_grad)
    return x


# ====================================================================== optim methods
class OptimMethod:
    """Base class. ``optimize(feval, x)`` evaluates ``feval(x) -> (loss, grad)`` and updates the flat vector ``x``
    in place, returning ``(x, [loss])``. ``update(x, grad)`` is the closure-free form the training loops use."""

    kind = -1

    def __init__(self):
        self.state = T()
        self.state["epoch"] = 1
        self.state["neval"] = 1
        self._hist: dict[str, torch.Tensor] = {}
        self.gradient_scale = 1.0        # extra factor folded into the fused update (1/replicas, clip, loss scale)
        self.clip_range: tuple[float, float] | None = None

    # ---- API parity
    def optimize(self, feval, parameter: torch.Tensor):
        fx, dfdx = feval(parameter)
        self.update(parameter, dfdx)
        return parameter, [fx]

    def update(self, x: torch.Tensor, g: torch.Tensor, w16: torch.Tensor | None = None, truncate: bool = False):
        raise NotImplementedError

    def native_plan(self, x: torch.Tensor, g: torch.Tensor | None = None):
        """Advance the per-step scalars (schedule, counters) and return ``(kind, hp_bytes, s0, s1)`` describing this
        step for the fused kernels (csrc/optim.cu, csrc/comm_fused.cu). ``x`` is the fp32 vector (or slice) that will
        be updated; state buffers are created lazily with its shape."""
        raise NotImplementedError(f"{type(self).__name__} has no fused-kernel form")

    def clearHistory(self):
        self._hist.clear()
        return self

    clear_history = clearHistory

    def getLearningRate(self) -> float:
        return getattr(self, "learningRate", 0.0)

    get_learning_rate = getLearningRate

    def updateHyperParameter(self):
        pass

    def getHyperParameter(self) -> str:
        return ""

    def loadFromTable(self, config: Table):
        for k, v in config.items():
            if isinstance(k, str) and hasattr(self, k):
                setattr(self, k, v)
        return self

    def clone(self):
        return copy.deepcopy(self)

    def _buf(self, name: str, like: torch.Tensor, fill: float = 0.0) -> torch.Tensor:
        b = self._hist.get(name)
        if b is None or b.shape != like.shape or b.device != like.device:
            b = torch.full_like(like, fill, dtype=torch.float32)
            self._hist[name] = b
        return b

    def _pre(self, g: torch.Tensor) -> torch.Tensor:
        """CPU path: apply gradient scale / constant clipping out of place."""
        if self.gradient_scale != 1.0:
            g = g * self.gradient_scale
        if self.clip_range is not None:
            g = g.clamp(self.clip_range[0], self.clip_range[1])
        return g

    def _hp_common(self):
        lo, hi = self.clip_range if self.clip_range is not None else (0.0, 0.0)
        return dict(gscale=self.gradient_scale, clip_min=lo, clip_max=hi)

    @staticmethod
    def _finish(x, w16, truncate):
        if w16 is not None:
            from ..ops.elementwise import cast_f32_to_bf16

            cast_f32_to_bf16(x, w16, truncate)

    # ---- persistence (S/optim/OptimMethod.scala:92-96: history is dropped on save)
    def save(self, path: str, overWrite: bool = False):
        from ..serialization.snapshot import save_optim_method

        save_optim_method(self, path, overWrite)
        return self

    @staticmethod
    def load(path: str):
        from ..serialization.snapshot import load_optim_method

        return load_optim_method(path)


class SGD(OptimMethod):
    kind = KIND_SGD

    def __init__(self, learningrate: float = 1e-3, learningrate_decay: float = 0.0, weightdecay: float = 0.0,
                 momentum: float = 0.0, dampening: float | None = None, nesterov: bool = False,
                 leaningrate_schedule: LearningRateSchedule | None = None, learningrates=None, weightdecays=None,
                 learningrate_schedule: LearningRateSchedule | None = None):
        super().__init__()
        self.learningRate, self.learningRateDecay, self.weightDecay = learningrate, learningrate_decay, weightdecay
        self.momentum = momentum
        self.dampening = momentum if dampening is None else dampening   # reference default: Double.MaxValue → momentum
        self.nesterov = nesterov
        self.learningRateSchedule = leaningrate_schedule or learningrate_schedule or Default()
        self.learningRates, self.weightDecays = learningrates, weightdecays
        if nesterov and (momentum <= 0 or self.dampening != 0):
            raise ValueError("Nesterov momentum requires a momentum and zero dampening")

    def updateHyperParameter(self):
        self.learningRateSchedule.updateHyperParameter(self)

    def getLearningRate(self):
        return self.learningRateSchedule.currentRate    # negative, like the reference

    def getHyperParameter(self):
        return f"Current learning rate is {-self.learningRateSchedule.currentRate}. "

    def _advance(self):
        self.learningRateSchedule.updateHyperParameter(self)
        first = ("v" not in self._hist) or bool(self.__dict__.pop("_hist_fresh", False))
        return -self.learningRateSchedule.currentRate, first

    def native_plan(self, x, g=None):
        if self.learningRates is not None or self.weightDecays is not None:
            raise NotImplementedError("per-parameter learningRates/weightDecays have no fused-kernel form")
        clr, first = self._advance()
        v = self._buf("v", x) if self.momentum != 0 else None
        hp = pack_hp(lr=clr, wd=self.weightDecay, momentum=self.momentum, dampening=self.dampening,
                     nesterov=self.nesterov, first_step=first and self.momentum != 0, **self._hp_common())
        return KIND_SGD, hp, v, None

    def update(self, x, g, w16=None, truncate=False):
        if x.is_cuda and self.learningRates is None and self.weightDecays is None:
            return run_plan(self.native_plan(x, g), x, g, w16, truncate)
        clr, first = self._advance()
        with torch.no_grad():
            g = self._pre(g)
            if self.weightDecays is not None:
                g = g + torch.as_tensor(self.weightDecays, dtype=x.dtype, device=x.device) * x
            elif self.weightDecay != 0:
                g = g + self.weightDecay * x
            if self.momentum != 0:
                if first:
                    v = self._buf("v", x)
                    v.copy_(g)
                else:
                    v = self._hist["v"]
                    v.mul_(self.momentum).add_(g, alpha=1 - self.dampening)
                g = g + self.momentum * v if self.nesterov else v
            if self.learningRates is not None:
                x.sub_(clr * torch.as_tensor(self.learningRates, dtype=x.dtype, device=x.device) * g)
            else:
                x.sub_(g, alpha=clr)
        self._finish(x, w16, truncate)
        return x


class Adam(OptimMethod):
    kind = KIND_ADAM

    def __init__(self, learningrate: float = 1e-3, learningrate_decay: float = 0.0, beta1: float = 0.9,
                 beta2: float = 0.999, epsilon: float = 1e-8):
        super().__init__()
        self.learningRate, self.learningRateDecay = learningrate, learningrate_decay
        self.beta1, self.beta2, self.Epsilon = beta1, beta2, epsilon

    def getHyperParameter(self):
        return f"Current learning rate is {self.learningRate}. "

    def _advance(self):
        t = self.state.getOrElse("evalCounter", 0)
        clr = self.learningRate / (1 + t * self.learningRateDecay)
        t += 1
        self.state["evalCounter"] = t
        return clr * math.sqrt(1 - self.beta2 ** t) / (1 - self.beta1 ** t)

    def native_plan(self, x, g=None):
        step = self._advance()
        hp = pack_hp(lr=step, momentum=self.beta1, dampening=self.beta2, eps=self.Epsilon, **self._hp_common())
        return KIND_ADAM, hp, self._buf("s", x), self._buf("r", x)

    def update(self, x, g, w16=None, truncate=False):
        if x.is_cuda:
            return run_plan(self.native_plan(x, g), x, g, w16, truncate)
        step = self._advance()
        m, v = self._buf("s", x), self._buf("r", x)
        with torch.no_grad():
            g = self._pre(g)
            m.mul_(self.beta1).add_(g, alpha=1 - self.beta1)
            v.mul_(self.beta2).addcmul_(g, g, value=1 - self.beta2)
            x.addcdiv_(m, v.sqrt().add_(self.Epsilon), value=-step)
        self._finish(x, w16, truncate)
        return x


class ParallelAdam(Adam):
    """Reference: chunked multi-thread Adam (S/optim/ParallelAdam.scala). One flat kernel already uses the whole GPU,
    so this is Adam with the ``parallel_num`` knob kept for API compatibility."""

    def __init__(self, learningrate=1e-3, learningrate_decay=0.0, beta1=0.9, beta2=0.999, epsilon=1e-8,
                 parallel_num: int = -1):
        super().__init__(learningrate, learningrate_decay, beta1, beta2, epsilon)
        self.parallel_num = parallel_num


class Adagrad(OptimMethod):
    kind = KIND_ADAGRAD

    def __init__(self, learningrate: float = 1e-3, learningrate_decay: float = 0.0, weightdecay: float = 0.0):
        super().__init__()
        self.learningRate, self.learningRateDecay, self.weightDecay = learningrate, learningrate_decay, weightdecay

    def _advance(self):
        n = self.state.getOrElse("evalCounter", 0)
        self.state["evalCounter"] = n + 1
        return self.learningRate / (1 + n * self.learningRateDecay)

    def native_plan(self, x, g=None):
        clr = self._advance()
        return KIND_ADAGRAD, pack_hp(lr=clr, wd=self.weightDecay, **self._hp_common()), self._buf("paramVariance", x), None

    def update(self, x, g, w16=None, truncate=False):
        if x.is_cuda:
            return run_plan(self.native_plan(x, g), x, g, w16, truncate)
        clr = self._advance()
        G = self._buf("paramVariance", x)
        with torch.no_grad():
            g = self._pre(g)
            if self.weightDecay != 0:
                g = g + self.weightDecay * x
            G.addcmul_(g, g)
            x.addcdiv_(g, G.sqrt().add_(1e-10), value=-clr)
        self._finish(x, w16, truncate)
        return x


class RMSprop(OptimMethod):
    kind = KIND_RMSPROP

    def __init__(self, learningrate: float = 1e-2, learningrate_decay: float = 0.0, decayrate: float = 0.99,
                 epsilon: float = 1e-8):
        super().__init__()
        self.learningRate, self.learningRateDecay, self.decayRate, self.Epsilon = (learningrate, learningrate_decay,
                                                                                 decayrate, epsilon)

    def _advance(self):
        n = self.state.getOrElse("evalCounter", 0)
        self.state["evalCounter"] = n + 1
        return self.learningRate / (1 + n * self.learningRateDecay)

    def native_plan(self, x, g=None):
        clr = self._advance()
        hp = pack_hp(lr=clr, momentum=self.decayRate, eps=self.Epsilon, **self._hp_common())
        return KIND_RMSPROP, hp, self._buf("sumSquare", x), None

    def update(self, x, g, w16=None, truncate=False):
        if x.is_cuda:
            return run_plan(self.native_plan(x, g), x, g, w16, truncate)
        clr = self._advance()
        ss = self._buf("sumSquare", x)
        with torch.no_grad():
            g = self._pre(g)
            ss.mul_(self.decayRate).addcmul_(g, g, value=1 - self.decayRate)
            x.addcdiv_(g, ss.sqrt().add_(self.Epsilon), value=-clr)
        self._finish(x, w16, truncate)
        return x


class Adadelta(OptimMethod):
    kind = KIND_ADADELTA

    def __init__(self, decayrate: float = 0.9, epsilon: float = 1e-10):
        super().__init__()
        self.decayRate, self.Epsilon = decayrate, epsilon

    def getLearningRate(self):
        return 0.0

    def _advance(self):
        self.state["evalCounter"] = self.state.getOrElse("evalCounter", 0) + 1

    def native_plan(self, x, g=None):
        self._advance()
        hp = pack_hp(momentum=self.decayRate, eps=self.Epsilon, **self._hp_common())
        return KIND_ADADELTA, hp, self._buf("paramVariance", x), self._buf("accDelta", x)

    def update(self, x, g, w16=None, truncate=False):
        if x.is_cuda:
            return run_plan(self.native_plan(x, g), x, g, w16, truncate)
        self._advance()
        var, acc = self._buf("paramVariance", x), self._buf("accDelta", x)
        with torch.no_grad():
            g = self._pre(g)
            var.mul_(self.decayRate).addcmul_(g, g, value=1 - self.decayRate)
            delta = (acc + self.Epsilon).sqrt_().div_((var + self.Epsilon).sqrt_()).mul_(g)
            x.sub_(delta)
            acc.mul_(self.decayRate).addcmul_(delta, delta, value=1 - self.decayRate)
        self._finish(x, w16, truncate)
        return x


class Adamax(OptimMethod):
    kind = KIND_ADAMAX

    def __init__(self, learningrate: float = 2e-3, beta1: float = 0.9, beta2: float = 0.999, epsilon: float = 1e-38):
        super().__init__()
        self.learningRate, self.beta1, self.beta2, self.Epsilon = learningrate, beta1, beta2, epsilon

    def _advance(self):
        t = self.state.getOrElse("evalCounter", 0) + 1
        self.state["evalCounter"] = t
        return self.learningRate / (1 - self.beta1 ** t)

    def native_plan(self, x, g=None):
        step = self._advance()
        hp = pack_hp(lr=step, momentum=self.beta1, dampening=self.beta2, eps=self.Epsilon, **self._hp_common())
        return KIND_ADAMAX, hp, self._buf("m", x), self._buf("u", x)

    def update(self, x, g, w16=None, truncate=False):
        if x.is_cuda:
            return run_plan(self.native_plan(x, g), x, g, w16, truncate)
        step = self._advance()
        m, u = self._buf("m", x), self._buf("u", x)
        with torch.no_grad():
            g = self._pre(g)
            m.mul_(self.beta1).add_(g, alpha=1 - self.beta1)
            torch.maximum(u * self.beta2, g.abs() + self.Epsilon, out=u)
            x.addcdiv_(m, u, value=-step)
        self._finish(x, w16, truncate)
        return x


class Ftrl(OptimMethod):
    kind = KIND_FTRL

    def __init__(self, learningrate: float = 1e-3, learningrate_power: float = -0.5,
                 initial_accumulator_value: float = 0.1, l1_regularization_strength: float = 0.0,
                 l2_regularization_strength: float = 0.0, l2_shrinkage_regularization_strength: float = 0.0):
        super().__init__()
        if learningrate < 0:
            raise ValueError(f"Ftrl: learning rate should be greater or equal to zero. but got {learningrate}")
        if learningrate_power > 0:
            raise ValueError("Ftrl: learning rate power should be smaller or equal to zero.")
        for nm, v in (("initial accumulator", initial_accumulator_value), ("L1", l1_regularization_strength),
                      ("L2", l2_regularization_strength), ("L2 shrinkage", l2_shrinkage_regularization_strength)):
            if v < 0:
                raise ValueError(f"Ftrl: {nm} value should be greater or equal to zero. but got {v}")
        self.learningRate, self.learningRatePower = learningrate, learningrate_power
        self.initialAccumulatorValue = initial_accumulator_value
        self.l1RegularizationStrength, self.l2RegularizationStrength = (l1_regularization_strength,
                                                                       l2_regularization_strength)
        self.l2ShrinkageRegularizationStrength = l2_shrinkage_regularization_strength

    def native_plan(self, x, g=None):
        hp = pack_hp(lr=self.learningRate, lr_power=self.learningRatePower, l1=self.l1RegularizationStrength,
                     l2=self.l2RegularizationStrength, l2_shrink=self.l2ShrinkageRegularizationStrength,
                     **self._hp_common())
        return KIND_FTRL, hp, self._buf("accum", x, self.initialAccumulatorValue), self._buf("linear", x)

    def update(self, x, g, w16=None, truncate=False):
        if x.is_cuda:
            return run_plan(self.native_plan(x, g), x, g, w16, truncate)
        accum = self._buf("accum", x, self.initialAccumulatorValue)
        linear = self._buf("linear", x)
        lr, lrp = self.learningRate, self.learningRatePower
        l1, l2, l2s = (self.l1RegularizationStrength, self.l2RegularizationStrength,
                       self.l2ShrinkageRegularizationStrength)
        with torch.no_grad():
            g = self._pre(g)
            gs = g + 2 * l2s * x
            accum_new = accum + gs * gs
            p_old, p_new = accum.pow(-lrp), accum_new.pow(-lrp)
            linear.add_(gs + (p_old - p_new) / lr * x)
            quadratic = p_new / lr + 2 * l2
            accum.copy_(accum_new)
            sgn = torch.sign(linear) * l1
            x.copy_(torch.where(linear.abs() > l1, (sgn - linear) / quadratic, torch.zeros_like(x)))
        self._finish(x, w16, truncate)
        return x


class LarsSGD(SGD):
    """Layer-wise adaptive rate scaling (S/optim/LarsSGD.scala:73-100). One instance per layer (ParallelOptimizer /
    per-submodule optim methods); the trust scale is either supplied through ``setGradientScale`` (computed globally
    by LarsProcessor) or derived from the local norms."""

    kind = KIND_LARS

    def __init__(self, lars_learning_rate_only: bool = False, trust: float = 1.0, learningrate: float = 1e-3,
                 learningrate_decay: float = 0.0, weightdecay: float = 0.0, momentum: float = 0.0,
                 learningrate_schedule: LearningRateSchedule | None = None):
        OptimMethod.__init__(self)
        if not (0.0 < trust <= 1.0):
            raise ValueError(f"the trust for LARS is {trust}, which should be greater than 0 and less than 1")
        self.larsLearningRateOnly, self.trust = lars_learning_rate_only, trust
        self.learningRate, self.learningRateDecay, self.weightDecay, self.momentum = (learningrate, learningrate_decay,
                                                                                    weightdecay, momentum)
        self.dampening, self.nesterov = 0.0, False
        self.learningRates = self.weightDecays = None
        self.learningRateSchedule = learningrate_schedule or Default()
        self._calculated_trust: float | None = None

    def setGradientScale(self, scale: float):
        self._calculated_trust = scale

    def _gradient_scale(self, g, x) -> float:
        if self._calculated_trust is not None:
            raw, self._calculated_trust = self._calculated_trust, None
        else:
            ng = float(g.double().pow(2).sum().sqrt()) * abs(self.gradient_scale)
            nx = float(x.double().pow(2).sum().sqrt())
            raw = (ng + self.weightDecay * nx) / nx if nx != 0 else float("inf")
        if math.isinf(raw):
            return 10000.0
        if math.isnan(raw):
            return 1.0
        if abs(raw) < 1e-4:
            return 1e-4
        return raw

    def _advance_lars(self, x, g):
        self.learningRateSchedule.updateHyperParameter(self)
        global_lr = -self.learningRateSchedule.currentRate * self.trust
        return global_lr / self._gradient_scale(g, x)

    def native_plan(self, x, g=None):
        rate = self._advance_lars(x, g)
        hp = pack_hp(lr=rate, wd=self.weightDecay, momentum=self.momentum, **self._hp_common())
        return KIND_LARS, hp, self._buf("v", x), None

    def update(self, x, g, w16=None, truncate=False):
        if x.is_cuda:
            return run_plan(self.native_plan(x, g), x, g, w16, truncate)
        rate = self._advance_lars(x, g)
        v = self._buf("v", x)
        with torch.no_grad():
            g = self._pre(g)
            v.mul_(self.momentum).add_((g + self.weightDecay * x) * rate)
            x.sub_(v)
        self._finish(x, w16, truncate)
        return x


class LBFGS(OptimMethod):
    """Limited-memory BFGS with optional line search (S/optim/LBFGS.scala:72-290, LineSearch.scala)."""

    def __init__(self, max_iter: int = 20, max_eval: float | None = None, tolfun: float = 1e-5, tolx: float = 1e-9,
                 ncorrection: int = 100, learningrate: float = 1.0, verbose: bool = False, linesearch=None,
                 linesearch_options=None):
        super().__init__()
        self.maxIter = max_iter
        self.maxEval = max_iter * 1.25 if max_eval is None else max_eval
        self.tolFun, self.tolX, self.nCorrection, self.learningRate = tolfun, tolx, ncorrection, learningrate
        self.verbose, self.lineSearch, self.lineSearchOptions = verbose, linesearch, linesearch_options

    def update(self, x, g, w16=None, truncate=False):
        raise RuntimeError("LBFGS needs the closure form: use optimize(feval, x)")

    def optimize(self, feval, x):
        st = self.state
        func_eval = st.getOrElse("funcEval", 0)
        n_iter = st.getOrElse("nIter", 0)
        f, g = feval(x)
        g = g.clone()
        f_hist = [f]
        current_evals = 1
        func_eval += 1
        if float(g.abs().sum()) <= self.tolFun:
            return x, f_hist
        d = st.get("dir")
        old_dirs = st.getOrElse("old_dirs", [])
        old_stps = st.getOrElse("old_stps", [])
        Hdiag = st.getOrElse("Hdiag", 1.0)
        g_old = st.get("g_old")
        f_old = st.getOrElse("f_old", 0.0)
        t = st.getOrElse("t", self.learningRate)
        it = 0
        while it < self.maxIter:
            it += 1
            n_iter += 1
            if n_iter == 1:
                d = -g
                old_dirs, old_stps, Hdiag = [], [], 1.0
            else:
                y = g - g_old
                s = d * t
                ys = float(y.dot(s))
                if ys > 1e-10:
                    if len(old_dirs) == self.nCorrection:
                        old_dirs.pop(0)
                        old_stps.pop(0)
                    old_dirs.append(s)
                    old_stps.append(y)
                    Hdiag = ys / float(y.dot(y))
                k = len(old_dirs)
                ro = [1.0 / float(old_stps[i].dot(old_dirs[i])) for i in range(k)]
                q = -g
                al = [0.0] * k
                for i in range(k - 1, -1, -1):
                    al[i] = float(old_dirs[i].dot(q)) * ro[i]
                    q = q - al[i] * old_stps[i]
                r = q * Hdiag
                for i in range(k):
                    be_i = float(old_stps[i].dot(r)) * ro[i]
                    r = r + (al[i] - be_i) * old_dirs[i]
                d = r
            g_old = g.clone()
            f_old = f
            gtd = float(g.dot(d))
            if gtd > -self.tolX:
                break
            t = min(1.0, 1.0 / float(g.abs().sum())) * self.learningRate if n_iter == 1 else self.learningRate
            ls_evals = 0
            if self.lineSearch is not None:
                f, g, x_new, t, ls_evals = self.lineSearch(feval, x, t, d, f, g, gtd, self.lineSearchOptions)
                x.copy_(x_new)
                f_hist.append(f)
            else:
                x.add_(d, alpha=t)
                if it != self.maxIter:
                    f, g = feval(x)
                    g = g.clone()
                    ls_evals = 1
                    f_hist.append(f)
            current_evals += ls_evals
            func_eval += ls_evals
            if it == self.maxIter or current_evals >= self.maxEval:
                break
            if float(g.abs().sum()) <= self.tolFun:
                break
            if float((d * t).abs().sum()) <= self.tolX:
                break
            if abs(f - f_old) < self.tolX:
                break
        st["old_dirs"], st["old_stps"], st["Hdiag"] = old_dirs, old_stps, Hdiag
        st["g_old"], st["f_old"], st["t"], st["dir"] = g_old, f_old, t, d
        st["funcEval"], st["nIter"] = func_eval, n_iter
        return x, f_hist

    def clearHistory(self):
        for k in ("old_dirs", "old_stps", "Hdiag", "g_old", "f_old", "t", "dir", "funcEval", "nIter"):
            self.state.delete(k)
        return self


class LineSearch:
    """Step-length search plugged into ``LBFGS(linesearch=...)`` (S/optim/LineSearch.scala:25-50).

    ``__call__(opfunc, x, t, d, f, g, gtd, options)`` gets the objective closure, the current point, the initial step
    ``t``, the direction ``d``, f(x), ∇f(x) and the directional derivative gᵀd, and returns
    ``(f_new, g_new, x_new, t, n_evals)``."""

    def __call__(self, opfunc, x, t, d, f, g, gtd, options=None):  # pragma: no cover - interface
        raise NotImplementedError


class WolfeLineSearch(LineSearch):
    """Strong-Wolfe line search: bracket a step that satisfies the sufficient-decrease (c1) and curvature (c2)
    conditions, then shrink the bracket by safeguarded cubic interpolation. The reference ships only the interface;
    this is the search Torch's ``optim.lbfgs`` users pass as ``lswolfe``. ``options``: c1, c2, tolX, maxIter."""

    def __init__(self, c1: float = 1e-4, c2: float = 0.9, tolX: float = 1e-9, maxIter: int = 20):
        self.c1, self.c2, self.tolX, self.maxIter = c1, c2, tolX, maxIter

    @staticmethod
    def _cubic_min(x1, f1, g1, x2, f2, g2, lo=None, hi=None):
        """Minimiser of the cubic through (x1, f1, g1), (x2, f2, g2), clamped to [lo, hi]; the midpoint when the
        cubic has no real minimiser."""
        if lo is None:
            lo, hi = (x1, x2) if x1 <= x2 else (x2, x1)
        d1 = g1 + g2 - 3 * (f1 - f2) / (x1 - x2)
        sq = d1 * d1 - g1 * g2
        if sq >= 0:
            d2 = sq ** 0.5
            if x1 <= x2:
                p = x2 - (x2 - x1) * ((g2 + d2 - d1) / (g2 - g1 + 2 * d2))
            else:
                p = x1 - (x1 - x2) * ((g1 + d2 - d1) / (g1 - g2 + 2 * d2))
            return min(max(p, lo), hi)
        return (lo + hi) / 2

    def __call__(self, opfunc, x, t, d, f, g, gtd, options=None):
        o = options or {}
        get = (lambda k, dflt: o.get(k, dflt)) if hasattr(o, "get") else (lambda k, dflt: dflt)
        c1, c2, tolx, max_iter = get("c1", self.c1), get("c2", self.c2), get("tolX", self.tolX), int(
            get("maxIter", self.maxIter))
        d_norm = float(d.abs().max())
        x0 = x.clone()

        def evaluate(step):
            fv, gv = opfunc(x0 + step * d)
            gv = gv.clone()
            return float(fv), gv, float(gv.dot(d))

        f0, g0, gtd0 = float(f), g.clone(), float(gtd)
        f_new, g_new, gtd_new = evaluate(t)
        evals = 1
        t_prev, f_prev, g_prev, gtd_prev = 0.0, f0, g0, gtd0
        bracket = None
        it = 0
        done = False
        while it < max_iter:
            if f_new > f0 + c1 * t * gtd0 or (it > 0 and f_new >= f_prev):
                bracket = [(t_prev, f_prev, g_prev, gtd_prev), (t, f_new, g_new, gtd_new)]
                break
            if abs(gtd_new) <= -c2 * gtd0:
                bracket, done = [(t, f_new, g_new, gtd_new)], True
                break
            if gtd_new >= 0:
                bracket = [(t_prev, f_prev, g_prev, gtd_prev), (t, f_new, g_new, gtd_new)]
                break
            lo, hi = t + 0.01 * (t - t_prev), t * 10
            nxt = self._cubic_min(t_prev, f_prev, gtd_prev, t, f_new, gtd_new, lo, hi)
            t_prev, f_prev, g_prev, gtd_prev = t, f_new, g_new, gtd_new
            t = nxt
            f_new, g_new, gtd_new = evaluate(t)
            evals += 1
            it += 1
        if bracket is None:                       # ran out of iterations while still descending
            bracket, done = [(t, f_new, g_new, gtd_new)], True

        stalled = False
        while not done and it < max_iter:
            (ta, fa, ga, gtda), (tb, fb, gb, gtdb) = bracket
            if abs(tb - ta) * d_norm < tolx:
                break
            t = self._cubic_min(ta, fa, gtda, tb, fb, gtdb)
            span, lo_b, hi_b = abs(tb - ta), min(ta, tb), max(ta, tb)
            # too close to an end of the bracket: nudge inwards (or bisect when that already happened once)
            if min(hi_b - t, t - lo_b) < 0.1 * span:
                if stalled or t >= hi_b or t <= lo_b:
                    t = hi_b - 0.1 * span if abs(t - hi_b) < abs(t - lo_b) else lo_b + 0.1 * span
                    stalled = False
                else:
                    stalled = True
            else:
                stalled = False
            f_new, g_new, gtd_new = evaluate(t)
            evals += 1
            it += 1
            lo_i = 0 if fa <= fb else 1
            t_lo, f_lo = bracket[lo_i][0], bracket[lo_i][1]
            if f_new > f0 + c1 * t * gtd0 or f_new >= f_lo:
                bracket[1 - lo_i] = (t, f_new, g_new, gtd_new)      # replaces the high end
            else:
                if abs(gtd_new) <= -c2 * gtd0:
                    done = True
                elif gtd_new * (bracket[1 - lo_i][0] - t_lo) >= 0:
                    bracket[1 - lo_i] = bracket[lo_i]
                bracket[lo_i] = (t, f_new, g_new, gtd_new)
        best = min(bracket, key=lambda b: b[1])
        t, f_new, g_new = best[0], best[1], best[2]
        return f_new, g_new, x0 + t * d, t, evals
