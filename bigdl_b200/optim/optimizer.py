"""Optimizer(model, dataset, criterion).optimize() — Local and Distri flavours.

Reference: S/optim/Optimizer.scala (builder API :111-463, factory :602-681), AbstractOptimizer.scala (validate
:93-191, checkpoint :205-231, saveSummary :47-79), LocalOptimizer.scala:95-214, DistriOptimizer.scala:97-517
(iteration), :803-983 (set-up, retry loop), Python surface P/optim/optimizer.py:724-1050.

B200 design: one process per GPU (launched with torchrun); "node" = rank. One iteration is
    H2D(batch) → forward → criterion → backward → AllReduceParameter.step (fused NVLink reduce-scatter + sharded
    OptimMethod update + all-gather) → bookkeeping
with no driver/executor split, so the two Spark jobs per iteration of the reference disappear. The iteration
accounting (``epoch``, ``neval``, ``recordsProcessedThisEpoch``, log line, triggers) follows the reference exactly.
"""
from __future__ import annotations

import logging
import os
import time
from datetime import datetime

import torch
import torch.distributed as dist

from ..dataset.base import (AbstractDataSet, DataSet, MiniBatch, Sample, SampleToMiniBatch)
from ..nn.criterion import Criterion
from ..nn.module import Module
from ..utils import config
from ..utils.engine import Engine
from ..utils.table import Table, T
from .metrics import Metrics
from .optim_method import SGD, OptimMethod
from .trigger import MaxEpoch, Trigger
from .validation import ValidationMethod, all_reduce_result

log = logging.getLogger("bigdl_b200.optim")


class GradientClipping:
    def __init__(self):
        self.constant: tuple[float, float] | None = None
        self.l2norm: float | None = None


class AbstractOptimizer:
    """What the local and the distributed optimizer share in the reference — checkpointing, validation and the end-of-
    training model hand-back (S/optim/AbstractOptimizer.scala:33-260). Here ``Optimizer`` carries that common code and
    ``_train_loop`` is the single loop both run; this is the common base name for isinstance checks."""


class Optimizer(AbstractOptimizer):
    """Builder + training loop base. Use ``Optimizer.create`` (or ``Optimizer(...)`` which dispatches the same way)."""

    def __new__(cls, *args, **kw):
        if cls is Optimizer:
            ds = kw.get("training_set", kw.get("dataset", args[1] if len(args) > 1 else None))
            distributed = (isinstance(ds, AbstractDataSet) and ds.is_distributed()) or \
                (dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1)
            return object.__new__(DistriOptimizer if distributed else LocalOptimizer)
        return object.__new__(cls)

    def __init__(self, model: Module, training_set=None, criterion: Criterion | None = None,
                 end_trigger: Trigger | None = None, batch_size: int | None = None,
                 optim_method: OptimMethod | dict | None = None, dataset=None, **_ignored):
        training_set = training_set if training_set is not None else dataset
        self.model = model
        self.criterion = criterion
        self.batch_size = batch_size
        self.dataset = self._as_dataset(training_set, batch_size)
        self.end_when = end_trigger or MaxEpoch(1)
        if isinstance(optim_method, dict):
            self.optim_methods = dict(optim_method)
        else:
            self.optim_methods = {model.getName(): optim_method or SGD()}
        self.state = T()
        self.validation_trigger = None
        self.validation_dataset = None
        self.validation_methods: list[ValidationMethod] = []
        self.checkpoint_path = None
        self.checkpoint_trigger = None
        self.overwrite_checkpoint = False
        self.train_summary = None
        self.validation_summary = None
        self.clipping = GradientClipping()
        self.drop_percentage = 0.0
        self.max_drop_percentage = 0.0
        self.compute_threshold_batch_size = 100
        self.warmup_iteration_num = 200
        self.metrics = Metrics()
        self.parameter_processors = []
        self._param_sync = None

    # ------------------------------------------------------------------ factory
    @staticmethod
    def create(model, training_set, criterion, end_trigger=None, batch_size=32, optim_method=None, cores=None,
               bigdl_type="float"):
        return Optimizer(model, training_set, criterion, end_trigger, batch_size, optim_method)

    @staticmethod
    def apply(model, dataset, criterion, batch_size=None, **kw):
        return Optimizer(model, dataset, criterion, batch_size=batch_size, **kw)

    def _as_dataset(self, ds, batch_size):
        if ds is None:
            return None
        if isinstance(ds, AbstractDataSet):
            first = None
            try:
                it = ds.data(train=False)
                first = next(iter(it))
            except StopIteration:
                pass
            if isinstance(first, Sample):
                if not batch_size:
                    raise ValueError("batch_size is required when the dataset yields Samples")
                world = dist.get_world_size() if (dist.is_initialized() and ds.is_distributed()) else 1
                return ds.transform(SampleToMiniBatch(max(batch_size // world, 1)))
            return ds
        if isinstance(ds, tuple) and len(ds) == 2:      # (X, Y) ndarrays → local dataset
            return self._as_dataset(DataSet.from_ndarrays(ds[0], ds[1]), batch_size)
        if isinstance(ds, (list,)):
            return self._as_dataset(DataSet.array(ds), batch_size)
        raise TypeError(f"unsupported training set type {type(ds)}")

    # ------------------------------------------------------------------ builder API (reference names + python names)
    def setOptimMethod(self, method: OptimMethod):
        self.optim_methods = {self.model.getName(): method}
        return self

    set_optim_method = setOptimMethod

    def setOptimMethods(self, methods: dict):
        self.optim_methods = dict(methods)
        return self

    set_optim_methods = setOptimMethods

    def setEndWhen(self, trigger: Trigger):
        self.end_when = trigger
        return self

    set_end_when = setEndWhen

    def setValidation(self, trigger=None, dataset=None, vMethods=None, batch_size=None, val_rdd=None, val_method=None):
        dataset = dataset if dataset is not None else val_rdd
        vMethods = vMethods if vMethods is not None else val_method
        self.validation_trigger = trigger
        self.validation_dataset = self._as_dataset(dataset, batch_size or self.batch_size)
        self.validation_methods = list(vMethods)
        return self

    set_validation = setValidation

    def setCheckpoint(self, path: str, trigger: Trigger, isOverWrite: bool = True):
        self.checkpoint_path, self.checkpoint_trigger = path, trigger
        self.overwrite_checkpoint = isOverWrite
        return self

    def set_checkpoint(self, checkpoint_trigger, checkpoint_path, isOverWrite=True):
        return self.setCheckpoint(checkpoint_path, checkpoint_trigger, isOverWrite)

    def overWriteCheckpoint(self):
        self.overwrite_checkpoint = True
        return self

    def getCheckpointPath(self):
        return self.checkpoint_path

    def setTrainSummary(self, summary):
        self.train_summary = summary
        return self

    set_train_summary = setTrainSummary

    def setValidationSummary(self, summary):
        self.validation_summary = summary
        return self

    set_val_summary = setValidationSummary

    def setState(self, state: Table):
        self.state = state
        return self

    def setModel(self, model):
        self.model = model
        return self

    set_model = setModel

    def setCriterion(self, criterion):
        self.criterion = criterion
        return self

    set_criterion = setCriterion

    def setTrainData(self, dataset, batch_size=None):
        self.dataset = self._as_dataset(dataset, batch_size or self.batch_size)
        return self

    set_traindata = setTrainData

    def setGradientClippingByl2Norm(self, l2NormThreshold: float):
        self.clipping.l2norm = float(l2NormThreshold)
        return self

    set_gradclip_l2norm = setGradientClippingByl2Norm

    def setConstantGradientClipping(self, min: float, max: float):
        if min > max:
            raise ValueError("min value can not be larger than max")
        self.clipping.constant = (float(min), float(max))
        return self

    set_gradclip_const = setConstantGradientClipping

    def disableGradientClipping(self):
        self.clipping = GradientClipping()
        return self

    disable_gradclip = disableGradientClipping

    def setDropModuleProperty(self, dropPercentage: float, maxDropPercentage: float, batchsize: int = 100,
                              warmupIteration: int = 200):
        """Straggler dropping (S/optim/Optimizer.scala:404-412, S/optim/DistriOptimizer.scala:241-244,342-344,423-449).
        A replica here is a rank: after ``warmupIteration`` iterations every rank's forward+backward time is compared
        with the k-th largest time of the last ``batchsize`` iterations; a rank above the threshold contributes no
        gradient to that iteration (the average runs over the ranks that were kept), and an iteration that lost more
        than ``maxDropPercentage`` of its ranks is discarded. The decision needs the ranks' times on the host every
        iteration, so a drop policy runs the step eagerly on the nccl / gloo parameter backends."""
        from ..parallel.straggler import DropPolicy

        self.drop_percentage, self.max_drop_percentage = dropPercentage, maxDropPercentage
        self.compute_threshold_batch_size, self.warmup_iteration_num = batchsize, warmupIteration
        self._drop_policy = DropPolicy(dropPercentage, maxDropPercentage, batchsize, warmupIteration) \
            if dropPercentage > 0 else None
        return self

    def prepareInput(self):
        return self

    def setIterationCallback(self, fn):
        """``fn(state, loss_tensor)`` runs at the end of every iteration (after the parameter update has been
        enqueued). Used by the perf harnesses to drop CUDA events without leaving the public ``optimize()`` path."""
        self._iteration_callback = fn
        return self

    prepare_input = prepareInput

    # ------------------------------------------------------------------ helpers shared by the loops
    def _single_method(self) -> OptimMethod | None:
        if len(self.optim_methods) == 1:
            return next(iter(self.optim_methods.values()))
        return None

    def _check_submodules(self, flat):
        """Per-submodule optim methods must tile the flat vector with contiguous, non-overlapping ranges
        (S/optim/Optimizer.scala:492-522)."""
        ranges = []
        for name in self.optim_methods:
            m = self.model if name == self.model.getName() else self.model.find(name)
            if m is None:
                raise ValueError(f"Optimizer: cannot find submodule {name}")
            off, ln = flat.offsets_of(m)
            ranges.append((off, ln, name))
        ranges.sort()
        pos = 0
        for off, ln, name in ranges:
            if off < pos:
                raise ValueError(f"Optimizer: parameters of submodule {name} overlap another submodule")
            pos = off + ln
        if sum(ln for _, ln, _ in ranges) < sum(n for _, _, _, n in flat.entries):
            raise ValueError("Optimizer: the optim methods do not cover every parameter of the model")
        return ranges

    def _to_device(self, batch: MiniBatch, device) -> MiniBatch:
        if device.type == "cuda":
            return batch.to(device, non_blocking=True)
        return batch

    def _header(self, state, iteration_records, wall, loss, throughput_time, epoch_size, records_this_epoch):
        return (f"[Epoch {state['epoch']} {records_this_epoch}/{epoch_size}][Iteration {state['neval']}]"
                f"[Wall Clock {wall:.6f}s] Trained {iteration_records} records in {throughput_time:.6f} seconds. "
                f"Throughput is {iteration_records / max(throughput_time, 1e-12):.4f} records/second. Loss is {loss}. ")

    def _validate(self, state):
        if self.validation_trigger is None or self.validation_dataset is None:
            return
        if not self.validation_trigger(state):
            return
        from .evaluator import run_validation

        results = run_validation(self.model, self.validation_dataset, self.validation_methods,
                                 distributed=self._is_distributed())
        for method, res in zip(self.validation_methods, results):
            log.info("%s is %s", method, res)
            score = res.result()[0]
            if self.validation_summary is not None:
                self.validation_summary.addScalar(str(method), score, state["neval"] - 1)
        if results:
            state["score"] = results[0].result()[0]
            for om in self.optim_methods.values():
                om.state["score"] = state["score"]
        self.model.training()

    def _is_distributed(self) -> bool:
        return dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1

    def _rank(self) -> int:
        return dist.get_rank() if self._is_distributed() else 0

    def _checkpoint(self, state, subdir: str | None, distri_names: bool):
        if self.checkpoint_path is None or self.checkpoint_trigger is None:
            return
        if not self.checkpoint_trigger(state):
            return
        self._gather_for_checkpoint()
        if self._rank() != 0:
            return
        path = os.path.join(self.checkpoint_path, subdir) if subdir else self.checkpoint_path
        os.makedirs(path, exist_ok=True)
        n = state["neval"]
        self.model.save(os.path.join(path, f"model.{n}"), self.overwrite_checkpoint)
        for name, om in self.optim_methods.items():
            om.state["epoch"], om.state["neval"] = state["epoch"], state["neval"]
            fname = f"optimMethod-{name}.{n}" if distri_names else f"state.{n}"
            om.save(os.path.join(path, fname), self.overwrite_checkpoint)

    def _gather_for_checkpoint(self):
        st = getattr(self, "_strategy", None)
        if st is not None:
            st.drain()
        if self._param_sync is not None:
            self._param_sync.gather_master_weights()

    def _summary(self, state, loss, throughput, lr):
        ts = self.train_summary
        if ts is None or self._rank() != 0:
            return
        it = state["neval"] - 1
        if ts.should_log("Loss", state):
            ts.addScalar("Loss", loss, it)
        if ts.should_log("Throughput", state):
            ts.addScalar("Throughput", throughput, it)
        if ts.should_log("LearningRate", state):
            ts.addScalar("LearningRate", lr, it)
        if ts.has_trigger("Parameters") and ts.should_log("Parameters", state):
            self._gather_for_checkpoint()
            for m in self.model.all_modules():
                for n, p in m._parameters.items():
                    ts.addHistogram(f"{m.getName()}/{n}", p.detach().float().cpu(), it)

    def optimize(self):  # pragma: no cover - implemented by subclasses
        raise NotImplementedError


class LocalOptimizer(Optimizer):
    """Single-process training (S/optim/LocalOptimizer.scala:95-214). The reference clones the model once per core and
    sums replica gradients; one GPU (or one CPU process with intra-op threads) needs a single replica."""

    def optimize(self):
        return _train_loop(self, distributed=False)


class DistriOptimizer(Optimizer):
    """Synchronous data-parallel training over ``torch.distributed`` ranks (S/optim/DistriOptimizer.scala)."""

    def optimize(self):
        retry_times = int(config.get("bigdl.failure.retryTimes"))
        retry_interval = float(config.get("bigdl.failure.retryTimeInterval"))
        retry_num, last_failure = 0, time.time()
        if self.checkpoint_path is not None:
            # one timestamped folder per optimize() call, fixed BEFORE the retry loop (DistriOptimizer.scala:874-880):
            # a retry must find the snapshots the failed attempt wrote
            self._checkpoint_dir = os.path.join(self.checkpoint_path, datetime.now().strftime("%Y%m%d_%H%M%S"))
        while True:
            try:
                return _train_loop(self, distributed=True)
            except (ValueError, TypeError):
                raise       # IllegalArgumentException analogue: configuration errors are never retried
            except Exception as e:  # noqa: BLE001 - mirror of the reference's catch-all retry (DistriOptimizer.scala:881-963)
                if self.checkpoint_path is None:
                    raise
                now = time.time()
                if now - last_failure > retry_times * retry_interval:
                    retry_num = 1           # failures far apart: the window resets
                else:
                    retry_num += 1
                last_failure = now
                if retry_num > retry_times:
                    raise
                log.error("Error: %s\nRetrying %d/%d from the latest snapshot", e, retry_num, retry_times)
                self._recover_from_checkpoint()

    def _recover_from_checkpoint(self):
        from ..serialization.snapshot import latest_file, load_optim_method, load_snapshot_into

        root = getattr(self, "_checkpoint_dir", None) or self.checkpoint_path
        # release the failed attempt's parameter plane first: on the p2p backend the model's bf16 compute weights
        # alias the symmetric buffer, so they are moved to private memory before that buffer is freed
        if self._param_sync is not None:
            self._param_sync.close()
            self._param_sync = None
        model_file = latest_file(root, "model") if os.path.isdir(root) else None
        if model_file is None:
            # no snapshot yet: retry from the model and optim methods as they are (DistriOptimizer.scala:925-938)
            log.warning("no model snapshot under %s; retrying from the current model", root)
            return
        load_snapshot_into(self.model, model_file)
        for name in list(self.optim_methods):
            f = latest_file(root, f"optimMethod-{name}")
            if f is not None:
                om = load_optim_method(f)
                om.clearHistory()
                self.optim_methods[name] = om


def _train_loop(opt: Optimizer, distributed: bool, strategy_factory=None):
    from ..parallel.all_reduce_parameter import AllReduceParameter

    model, criterion, dataset = opt.model, opt.criterion, opt.dataset
    Engine.ensure()
    if not Engine.checkSingleton():
        raise RuntimeError("Detect multi-task run on one Executor/Container (bigdl.check.singleton)")
    model.checkDuplicate()
    model.training()
    if next(iter(model.torch_parameters()), torch.zeros(0)).is_cuda:
        from ..nn.criterion import fuse_final_logsoftmax

        fuse_final_logsoftmax(model, criterion)      # LogSoftMax + ClassNLL -> fused cross-entropy kernels (training only)
    device = model.device()
    weights, grads = model.getParameters()
    flat = model._flat
    world = dist.get_world_size() if (distributed and dist.is_initialized()) else 1
    ranges = opt._check_submodules(flat) if len(opt.optim_methods) > 1 else None

    prev_sync = model.__dict__.pop("_param_sync_keepalive", None)
    if prev_sync is not None:
        prev_sync.close()       # an earlier optimize() parked its symmetric buffer on the model; release it (collective)
    # ParallelOptimizer keeps full (non-sharded) weights and optimizer state on every rank: no sharded master copy
    method_ranges = [(o, l) for o, l, _ in ranges] if ranges is not None else None
    method_list = [opt.optim_methods[n] for _, _, n in ranges] if ranges is not None else None
    drop = getattr(opt, "_drop_policy", None) if (world > 1 and strategy_factory is None) else None
    backend = None if (world > 1 and strategy_factory is None) else "local"
    if drop is not None and device.type == "cuda" and config.get("bigdl.parameter.backend") in ("auto", "p2p"):
        backend = "nccl"         # per-iteration host decisions (who is dropped) do not fit the captured p2p step
    sync = AllReduceParameter(flat, backend=backend, root=model, method_ranges=method_ranges)
    sync.init()
    opt._param_sync = sync
    single = opt._single_method()
    if single is None and not sync.overlapped and sync.backend not in ("local", "gloo", "nccl"):
        raise NotImplementedError("per-submodule optim methods run on the local / gloo / bucketed-p2p parameter backends")

    for om in opt.optim_methods.values():
        om.clearHistory()        # DistriOptimizer.scala:811 — history never survives into a new optimize() call
        for k in ("epoch", "neval"):
            if k in opt.state:
                om.state[k] = opt.state[k]
    lead = next(iter(opt.optim_methods.values()))
    state = lead.state
    state["epoch"] = state.getOrElse("epoch", 1)
    state["neval"] = state.getOrElse("neval", 1)
    records_this_epoch = state.getOrElse("recordsProcessedThisEpoch", 0)
    epoch_size = dataset.size()
    if opt.checkpoint_path is not None:
        if not distributed:
            opt._checkpoint_dir = opt.checkpoint_path
        elif getattr(opt, "_checkpoint_dir", None) is None:
            opt._checkpoint_dir = os.path.join(opt.checkpoint_path, datetime.now().strftime("%Y%m%d_%H%M%S"))
    dataset.shuffle()
    data_iter = dataset.data(train=True)
    prefetch = _Prefetcher(data_iter, device) if device.type == "cuda" else None
    wall_start = time.time()
    clip = opt.clipping
    pending_loss = None        # GPU: the loss is read back one iteration late so the host never stalls the stream
    loss_value = float("nan")
    use_fused_loss = device.type == "cuda"
    graphs: dict = {}
    # per-submodule methods and global-norm clipping stay on the graph when the bucketed p2p plane carries them (device
    # resident OptimHP table and clip scale); on one GPU they run eagerly
    graph_ok = (device.type == "cuda" and bool(config.get("bigdl.cudaGraph")) and sync.backend in ("local", "p2p")
                and drop is None
                and config.get("bigdl.engineType") == "b200"
                and ((single is not None and clip.l2norm is None) or sync.overlapped))
    step_methods = method_list if method_list is not None else [single]
    strategy = strategy_factory(model, world) if strategy_factory is not None else None
    if strategy is not None:
        graph_ok = False
        opt._strategy = strategy

    # phase timers of the reference (S/optim/DistriOptimizer.scala:188-196): host clocks, so they are only taken when asked
    # for (bigdl.metrics) — they need a stream synchronisation per phase — or when a drop policy needs compute times
    metrics_on = bool(config.get("bigdl.metrics"))
    for name in ("get weights average", "computing time average", "aggregate gradient time", "data fetch time"):
        opt.metrics.set(name, 0.0, 1)
    while not opt.end_when(state):
        t0 = time.time()
        batch = next(prefetch) if prefetch is not None else next(data_iter)
        opt.metrics.add("data fetch time", (time.time() - t0) * 1e9)
        if not isinstance(batch, MiniBatch):
            raise TypeError("the training dataset must yield MiniBatch (use SampleToMiniBatch or pass batch_size)")
        n_local = batch.size()
        for om in opt.optim_methods.values():
            om.clip_range = clip.constant
        loss = None
        if graph_ok and len(batch.inputs) == 1 and len(batch.targets) == 1:
            loss = _graph_step(opt, graphs, model, criterion, sync, step_methods if sync.overlapped else single, batch,
                               device, world, clip.l2norm)
            if loss is None:
                graph_ok = False
        if loss is None:
            batch = opt._to_device(batch, device)
            x, y = batch.getInput(), batch.getTarget()
            if sync.overlapped:
                for om_ in step_methods:
                    om_.clip_range = clip.constant
                sync.begin_step(step_methods, world, clip.l2norm)
            t_compute = time.perf_counter()
            with torch.enable_grad():
                out = model._call(x)
                loss = criterion._call(out, y)
            seed = torch.ones((), dtype=loss.dtype, device=loss.device)
            seed._bdl_is_one = True
            torch.autograd.backward(loss, seed)
            if device.type == "cuda":
                from ..ops.functions import join_side_streams

                join_side_streams()
            model._apply_regularizers()
            if device.type == "cuda" and (drop is not None or metrics_on):
                torch.cuda.current_stream(device).synchronize()       # the phase timers below are host clocks
            compute_s = time.perf_counter() - t_compute
            opt.metrics.add("computing time average", compute_s * 1e9)
            t_agg = time.perf_counter()
            kept = world
            if drop is not None:
                kept, discard = _apply_drop_policy(drop, state["neval"], compute_s, flat, world, device)
                opt.metrics.set("dropped replicas", drop.dropped_total)
                if discard:            # too many ranks were late: nobody updates (DistriOptimizer.scala:342-344)
                    flat.grads.zero_()
                    kept = 0
            if kept == 0:
                pass
            elif strategy is not None:
                for om_ in strategy.methods:
                    om_.clip_range = clip.constant
                strategy.after_backward()
            elif sync.overlapped:
                sync.end_step()
            elif single is not None:
                sync.step(single, avg_numbers=kept, l2_clip=clip.l2norm)
            else:
                _multi_method_step(opt, sync, flat, ranges, world, clip)
            if metrics_on:
                if device.type == "cuda":
                    torch.cuda.current_stream(device).synchronize()
                opt.metrics.add("aggregate gradient time", (time.perf_counter() - t_agg) * 1e9)

        # ---- bookkeeping (DistriOptimizer.scala:399-474)
        if device.type == "cuda":
            # D2H read of the step's result, lagging one iteration so the host never stalls the stream it feeds
            snap = torch.empty(1, dtype=torch.float32).pin_memory() if not hasattr(opt, "_loss_ring") else None
            if snap is not None:
                opt._loss_ring = [snap, torch.empty(1, dtype=torch.float32).pin_memory()]
                opt._loss_events = [torch.cuda.Event(), torch.cuda.Event()]
                opt._loss_slot = 0
            slot = opt._loss_slot
            opt._loss_ring[slot].copy_(loss.detach().reshape(1).float(), non_blocking=True)
            opt._loss_events[slot].record()
            prev, pending_loss = pending_loss, (slot,)
            opt._loss_slot = 1 - slot
            if prev is not None:
                opt._loss_events[prev[0]].synchronize()
                loss_value = float(opt._loss_ring[prev[0]][0])

        else:
            loss_value = float(loss.detach())
        if distributed and world > 1 and device.type != "cuda":
            t = torch.tensor([loss_value], dtype=torch.float64)
            dist.all_reduce(t)
            loss_value = float(t) / world
        n_global = n_local * world
        records_this_epoch += n_global
        dt = time.time() - t0
        state["Loss"] = loss_value
        for om in opt.optim_methods.values():
            om.state["Loss"] = loss_value
        lr = lead.getLearningRate()
        if opt._rank() == 0:
            log.info(opt._header(state, n_global, time.time() - wall_start, loss_value, dt, epoch_size,
                                 records_this_epoch) + lead.getHyperParameter())
        state["neval"] = state["neval"] + 1
        state["recordsProcessedThisEpoch"] = records_this_epoch
        if records_this_epoch >= epoch_size:
            state["epoch"] = state["epoch"] + 1
            dataset.shuffle()
            data_iter = dataset.data(train=True)
            if prefetch is not None:
                prefetch.reset(data_iter)
            records_this_epoch = 0
            state["recordsProcessedThisEpoch"] = 0
        for om in opt.optim_methods.values():
            if om is not lead:
                om.state["epoch"], om.state["neval"] = state["epoch"], state["neval"]
        if strategy is not None:
            strategy.sync_states(state)
        cb = getattr(opt, "_iteration_callback", None)
        if cb is not None:
            cb(state, loss)
        opt._summary(state, loss_value, n_global / max(dt, 1e-12), lr)
        opt._validate(state)
        opt._checkpoint(state, os.path.basename(opt._checkpoint_dir) if (distributed and opt.checkpoint_path) else None,
                        distri_names=distributed)

    if pending_loss is not None:
        opt._loss_events[pending_loss[0]].synchronize()
        state["Loss"] = float(opt._loss_ring[pending_loss[0]][0])
    if strategy is not None:
        strategy.drain()           # apply the last iteration's averaged gradients
        strategy.detach()
    sync.gather_master_weights()
    if flat.compute is not None and sync.backend == "local":
        flat.dirty = True
    opt.state = state
    if metrics_on and opt._rank() == 0:
        log.info(opt.metrics.summary())
    Engine.resetSingleton()
    if sync.backend == "p2p":
        # keep the symmetric buffers alive while the model still points its compute weights at them
        model._param_sync_keepalive = sync
    return model


class _Prefetcher:
    """Pulls MiniBatches one step ahead and uploads them on a dedicated copy stream, so the host→device transfer of
    batch k+1 overlaps the compute of batch k (the GPU analogue of the reference's multi-threaded batchers,
    S/dataset/image/MTLabeledBGRImgToBatch.scala)."""

    def __init__(self, it, device):
        self.it, self.device = it, device
        self.stream = torch.cuda.Stream(device)
        self._next = None
        self._preload()

    def _preload(self):
        try:
            b = next(self.it)
        except StopIteration:
            self._next = None
            return
        if not isinstance(b, MiniBatch):
            self._next = (b, None)
            return
        with torch.cuda.stream(self.stream):
            dev = MiniBatch([t.to(self.device, non_blocking=True) for t in b.inputs],
                            [t.to(self.device, non_blocking=True) for t in b.targets])
            ev = torch.cuda.Event()
            ev.record(self.stream)
        self._next = (dev, ev)

    def reset(self, it):
        self.it = it
        self._preload()

    def __next__(self):
        if self._next is None:
            raise StopIteration
        b, ev = self._next
        if ev is not None:
            torch.cuda.current_stream(self.device).wait_event(ev)
            for t in b.inputs + b.targets:
                t.record_stream(torch.cuda.current_stream(self.device))
        self._preload()
        return b


def _apply_drop_policy(policy, iteration, compute_s, flat, world, device):
    """Exchange the ranks' compute times, update the threshold, and zero THIS rank's gradients when it was too slow.
    Returns (number of ranks kept, discard the iteration)."""
    t = torch.tensor([compute_s], dtype=torch.float64, device=device if device.type == "cuda" else "cpu")
    every = [torch.zeros_like(t) for _ in range(world)]
    dist.all_gather(every, t)
    times = [float(v) for v in every]
    keep, discard = policy.apply(times)
    policy.record(iteration, times)
    if dist.get_rank() not in keep:
        flat.grads.zero_()
    return len(keep), discard


def _graph_step(opt, graphs, model, criterion, sync, om, batch, device, world, l2_clip=None):
    """Run one iteration through a captured CUDA graph (one graph per batch shape). Returns the loss tensor, or None
    when this configuration cannot be captured (the caller then falls back to eager execution for good)."""
    from .graph_step import GraphedTrainStep

    x, y = batch.inputs[0], batch.targets[0]
    key = (tuple(x.shape), x.dtype, tuple(y.shape), y.dtype)
    gs = graphs.get(key)
    if gs is None:
        try:
            gs = GraphedTrainStep(model, criterion, sync, om, x.to(device), y.to(device), world, l2_clip=l2_clip)
        except NotImplementedError as e:
            log.warning("CUDA-graph capture unavailable for this configuration (%s); running eagerly", e)
            return None
        graphs[key] = gs        # capture warm-ups were rolled back; fall through to the first real replay
    return gs.run(x, y)


def _multi_method_step(opt, sync, flat, ranges, world, clip):
    """Per-submodule OptimMethods on the sharded vector (S/optim/DistriOptimizer.scala:370-389): every method updates
    the intersection of its parameter range with this rank's slice."""
    import torch.distributed as dist_

    g = flat.grads
    if world > 1:
        dist_.all_reduce(g)
    g.div_(world)
    scale = 1.0
    if clip.l2norm is not None:
        nrm = float(g.double().pow(2).sum().sqrt())
        scale = clip.l2norm / nrm if nrm > clip.l2norm else 1.0
    for off, ln, name in ranges:
        om = opt.optim_methods[name]
        om.gradient_scale = scale
        if ln:
            om.update(flat.weights[off:off + ln], g[off:off + ln],
                      w16=None if flat.compute is None else flat.compute[off:off + ln])
    g.zero_()
    flat.dirty = flat.compute is None
