"""P/optim/optimizer.py surface: optim methods, LR schedules, triggers, validation methods, Optimizer factories,
summaries, regularizers."""
from bigdl_b200.optim.optim_method import (SGD, Adadelta, Adagrad, Adam, Adamax, Default, EpochDecay,  # noqa: F401
                                           EpochDecayWithWarmUp, EpochSchedule, EpochStep, Exponential, Ftrl, LarsSGD,
                                           LBFGS, MultiStep, NaturalExp, OptimMethod, ParallelAdam, Plateau, Poly, Regime,
                                           RMSprop, SequentialSchedule, Step, Warmup)
from bigdl_b200.optim.optimizer import DistriOptimizer, LocalOptimizer, Optimizer  # noqa: F401
from bigdl_b200.optim.parallel_optimizer import ParallelOptimizer  # noqa: F401
from bigdl_b200.optim.regularizer import L1L2Regularizer, L1Regularizer, L2Regularizer  # noqa: F401
from bigdl_b200.optim.trigger import (EveryEpoch, MaxEpoch, MaxIteration, MaxScore, MinLoss, SeveralIteration,  # noqa: F401
                                      Trigger, TriggerAnd, TriggerOr)
from bigdl_b200.optim.validation import (HitRatio, Loss, MAE, MeanAveragePrecision, NDCG, Top1Accuracy,  # noqa: F401
                                         Top5Accuracy, TreeNNAccuracy)
from bigdl_b200.visualization import TrainSummary, ValidationSummary  # noqa: F401

BaseOptimizer = Optimizer


class ActivityRegularization:          # kept for import parity; the layer lives in bigdl.nn.layer
    pass

from bigdl_b200.optim.validation import MeanAveragePrecisionObjectDetection  # noqa: F401,E402
