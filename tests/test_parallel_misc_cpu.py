"""Parameter-plane utilities: truncated-bf16 codec classes (T/parameters/FP16ParameterSpec.scala vectors), clipping /
LARS processors, straggler threshold, thread ParameterSynchronizer, ModelBroadcast over gloo."""
import threading

import numpy as np
import pytest
import torch

from dist_utils import run_ranks


def test_fp16_compressed_tensor_reference_vectors():
    from bigdl_b200.parallel import FP16CompressedTensor, FP16SplitsCompressedTensor, SerializerInstance

    t = torch.tensor([1.111111, 4.111111, -1.111111, 0.0])
    c = SerializerInstance.create(t, "fp16")
    out = torch.zeros(4)
    c.deCompress(out)
    assert out.tolist() == [1.109375, 4.09375, -1.109375, 0.0]       # FP16ParameterSpec.scala:47-66
    # add = expand, add, truncate again
    c2 = FP16CompressedTensor(torch.tensor([2.222222, 1.0, 1.0, 1.0]))
    c.add(c2)
    c.deCompress(out)
    a, b = torch.tensor([1.109375, 4.09375, -1.109375, 0.0]), torch.zeros(4)
    c2.deCompress(b)
    want = ((a + b).view(torch.int32) >> 16 << 16).view(torch.float32)
    assert torch.equal(out, want)
    # byte export is big-endian upper halves of the floats
    raw = FP16CompressedTensor(torch.tensor([1.0])).bytes()
    assert raw == bytes([0x3F, 0x80])
    assert torch.equal(FP16CompressedTensor(raw).buf, FP16CompressedTensor(torch.tensor([1.0])).buf)
    # large buffers go through the threaded C++ codec and agree with the bit formula
    big = torch.randn(200000)
    cb = FP16CompressedTensor(big)
    assert torch.equal(cb.buf, (big.view(torch.int32) >> 16).to(torch.int16))
    other = FP16CompressedTensor(torch.randn(200000))
    ref = ((cb.buf.to(torch.int32) << 16).view(torch.float32) + (other.buf.to(torch.int32) << 16).view(torch.float32))
    cb.parAdd(other)
    assert torch.equal(cb.buf, (ref.view(torch.int32) >> 16).to(torch.int16))
    with pytest.raises(ValueError):
        SerializerInstance.create(t, "fp32")
    sp = FP16SplitsCompressedTensor(10, 3)
    src = torch.arange(10, dtype=torch.float32) + 0.111111
    sp.compress(src)
    back = torch.zeros(10)
    sp.deCompress(back)
    assert torch.equal(back, ((src.view(torch.int32) >> 16) << 16).view(torch.float32))
    with pytest.raises(NotImplementedError):
        sp.add(sp)


def test_clipping_processors_single_rank():
    from bigdl_b200.parallel import ConstantClippingProcessor, L2NormClippingProcessor

    g = torch.tensor([-3.0, 0.5, 4.0])
    ConstantClippingProcessor(-1, 1).processParameters(g, {})
    assert g.tolist() == [-1.0, 0.5, 1.0]
    g = torch.tensor([3.0, 4.0])
    st = {}
    p = L2NormClippingProcessor(1.0)
    p.collectGlobalData(g, st)
    p.processParameters(g, st)
    assert abs(st["l2Norm"] - 5.0) < 1e-6 and torch.allclose(g, torch.tensor([0.6, 0.8]))


def test_lars_processor_sets_trust_ratio():
    from bigdl_b200.optim.optim_method import LarsSGD
    from bigdl_b200.parallel import LarsProcessor

    w = torch.tensor([3.0, 4.0, 1.0, 0.0])
    g = torch.tensor([0.6, 0.8, 0.0, 2.0])
    a, b = LarsSGD(weightdecay=0.1, learningrate=1.0), LarsSGD(weightdecay=0.0, learningrate=1.0)
    proc = LarsProcessor([(0, 2, a), (2, 2, b)])
    st = {}
    proc.collectGlobalData(g, st, weight_slice=w)
    proc.processParameters(g, st)
    assert abs(a._calculated_trust - (1.0 + 0.1 * 5.0) / 5.0) < 1e-6
    assert abs(b._calculated_trust - 2.0) < 1e-6


def test_kth_largest_and_drop_policy():
    from bigdl_b200.parallel import DropPolicy, kth_largest

    vals = [5.0, 1.0, 9.0, 3.0, 7.0]
    assert [kth_largest(vals, k) for k in range(1, 6)] == [9.0, 7.0, 5.0, 3.0, 1.0]
    assert kth_largest(vals, 0) == float("inf")
    pol = DropPolicy(0.25, 0.5, batch_size=2, warmup_iteration=1)
    pol.record(1, [1, 1, 1, 1])                    # warm-up: ignored
    assert pol.threshold == float("inf")
    pol.record(2, [1.0, 1.1, 1.2, 5.0])
    pol.record(3, [1.0, 1.1, 1.2, 6.0])            # window full: k = 2 → threshold = 2nd largest = 5.0
    assert pol.threshold == 5.0
    keep, discard = pol.apply([1.0, 1.0, 1.0, 7.0])
    assert keep == [0, 1, 2] and not discard
    keep, discard = pol.apply([9.0, 9.0, 9.0, 1.0])
    assert keep == [3] and discard


def test_thread_parameter_synchronizer_sums_across_replicas():
    from bigdl_b200.parallel import ParameterSynchronizer

    ParameterSynchronizer.reset()
    ParameterSynchronizer.register("bn1", 4)
    out = [None] * 4

    def replica(i):
        ParameterSynchronizer.syncData("bn1", torch.full((3,), float(i + 1)), replica_id=i)
        out[i] = ParameterSynchronizer.collect("bn1")

    ts = [threading.Thread(target=replica, args=(i,)) for i in range(4)]
    [t.start() for t in ts]
    [t.join(30) for t in ts]
    assert all(torch.equal(o, torch.full((3,), 10.0)) for o in out)


def _bcast_worker(rank, world):
    from bigdl_b200.nn import layer as L
    from bigdl_b200.parallel import ModelBroadcast
    from bigdl_b200.utils.random_generator import RNG

    model = None
    if rank == 0:
        RNG.setSeed(77)
        model = L.Sequential().add(L.Linear(5, 7)).add(L.BatchNormalization(7)).add(L.ReLU()).add(L.Linear(7, 3))
        model.modules[1]._buffers["runningMean"].fill_(0.25)
    m = ModelBroadcast().broadcast(model).value(initGradient=True)
    w, g = m.getParameters()
    return w.clone().numpy(), float(m.modules[1]._buffers["runningMean"][0]), [type(c).__name__ for c in m.modules]


def test_model_broadcast_gloo():
    res = run_ranks(_bcast_worker, 2)
    assert np.array_equal(res[0][0], res[1][0]) and res[0][0].size > 0
    assert res[1][1] == 0.25
    assert res[0][2] == res[1][2] == ["Linear", "BatchNormalization", "ReLU", "Linear"]
