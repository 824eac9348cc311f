"""Shape-level checks of the detection heads, the less common recurrent containers and a few TF helper ops that no other
test instantiates (S/nn/{Proposal,RegionProposal,DetectionOutputSSD,DetectionOutputFrcnn,RecurrentDecoder,MultiRNNCell,
ConvLSTMPeephole3D}.scala, S/nn/tf/*)."""
import numpy as np
import torch

import bigdl_b200.nn.tf as TF
from bigdl_b200.nn import layer as L
from bigdl_b200.utils.random_generator import RNG
from bigdl_b200.utils.table import T


def test_recurrent_variants():
    RNG.setSeed(4)
    dec = L.RecurrentDecoder(4).add(L.LSTM(5, 5))
    y = dec.forward(torch.randn(2, 5))
    assert tuple(y.shape) == (2, 4, 5)                     # feeds its own output back for 4 steps
    assert tuple(dec.backward(torch.randn(2, 5), torch.ones_like(y)).shape) == (2, 5)
    stack = L.Recurrent().add(L.MultiRNNCell([L.LSTM(4, 6), L.LSTM(6, 3)]))
    x = torch.randn(2, 5, 4)
    assert tuple(stack.forward(x).shape) == (2, 5, 3)
    assert tuple(stack.backward(x, torch.ones(2, 5, 3)).shape) == (2, 5, 4)
    conv3d = L.Recurrent().add(L.ConvLSTMPeephole3D(2, 3, 3, 3))
    assert tuple(conv3d.forward(torch.randn(2, 3, 2, 4, 4, 4)).shape) == (2, 3, 3, 4, 4, 4)


def test_detection_heads_produce_bounded_boxes():
    RNG.setSeed(5)
    torch.manual_seed(5)
    rpn = L.RegionProposal(8, [32, 64], [0.5, 1.0, 2.0], [8, 16], 50, 20)
    rpn.evaluate()
    boxes = rpn.forward(T(T(torch.randn(1, 8, 16, 16), torch.randn(1, 8, 8, 8)), torch.tensor([[128., 128.]])))
    assert boxes.dim() == 2 and boxes.shape[1] == 4 and 0 < boxes.shape[0] <= 20
    assert float(boxes.min()) >= 0.0 and float(boxes.max()) <= 128.0
    prop = L.Proposal(100, 20, [0.5, 1.0, 2.0], [8.0, 16.0])
    prop.evaluate()
    A = 6
    rois = prop.forward(T(torch.rand(1, 2 * A, 6, 6), torch.randn(1, 4 * A, 6, 6) * 0.1, torch.tensor([[96., 96., 1., 1.]])))
    assert rois.shape[1] == 5 and 0 < rois.shape[0] <= 20 and float(rois[:, 1:].min()) >= 0.0
    frcnn = L.DetectionOutputFrcnn(0.3, 4, False, 10, 0.05)
    frcnn.evaluate()
    r = torch.tensor([[0., 10, 10, 50, 50], [0, 20, 20, 80, 80], [0, 5, 5, 30, 30]])
    out = frcnn.forward(T(torch.tensor([[96., 96., 1., 1.]]), r, torch.softmax(torch.randn(3, 4), 1), torch.randn(3, 16) * 0.1))
    n = int(out.reshape(-1)[0])
    assert out.shape[0] == 1 and out.numel() == 1 + 6 * n         # [count, (label, score, x1, y1, x2, y2) * count]
    P = 8
    ssd = L.DetectionOutputSSD(3, keep_top_k=5, nms_topk=10)
    ssd.evaluate()
    det = ssd.forward(T(torch.randn(1, P * 4) * 0.1, torch.randn(1, P * 3), torch.rand(1, 2, P * 4)))
    assert det.shape == (1, 1 + 6 * 5) and 0 <= int(det[0, 0]) <= 5


def test_tf_helper_ops():
    import cv2

    assert torch.allclose(TF.Log1p().forward(torch.tensor([0.5, 1.0])), torch.log1p(torch.tensor([0.5, 1.0])))
    assert torch.equal(TF.NoOp().forward(torch.tensor([1.0])), torch.tensor([1.0]))
    assert torch.equal(TF.Variable(torch.ones(2, 2)).forward(torch.zeros(1)), torch.ones(2, 2))
    offs = TF.ConcatOffset().forward(T(torch.tensor(1), torch.tensor([2, 3]), torch.tensor([2, 5])))
    assert [o.tolist() for o in offs.to_list()] == [[0, 0], [0, 3]]
    ok, buf = cv2.imencode(".png", np.random.RandomState(0).randint(0, 255, (6, 7, 3)).astype(np.uint8))
    assert ok and tuple(TF.DecodeImage(3).forward(buf.tobytes()).shape) == (6, 7, 3)


def test_tensor_array_stack_and_assign_ops():
    """TF data-flow helpers (S/nn/tf/{DataFlowOps,StateOps}.scala): a TensorArray written, scattered, read, concatenated
    and split; a stack pushed and popped; Assign copying into a Variable's storage."""
    handle, flow = TF.TensorArrayCreator().forward(torch.tensor(3)).to_list()
    TF.TensorArrayWrite().forward(T(handle, torch.tensor(0), torch.tensor([1., 2.]), flow))
    TF.TensorArrayScatter().forward(T(handle, torch.tensor([1, 2]), torch.tensor([[3., 4.], [5., 6.]]), flow))
    assert torch.equal(TF.TensorArrayRead().forward(T(handle, torch.tensor(2), flow)), torch.tensor([5., 6.]))
    assert int(TF.TensorArraySize().forward(T(handle, flow))) == 3
    assert torch.equal(TF.TensorArrayGather().forward(T(handle, torch.tensor([2, 0]), flow)),
                       torch.tensor([[5., 6.], [1., 2.]]))
    cat, lens = TF.TensorArrayConcat().forward(T(handle, flow)).to_list()
    assert torch.equal(cat, torch.tensor([1., 2., 3., 4., 5., 6.])) and lens.tolist() == [2, 2, 2]
    h2, f2 = TF.TensorArrayCreator().forward(torch.tensor(2)).to_list()
    TF.TensorArraySplit().forward(T(h2, torch.arange(5.), torch.tensor([2, 3]), f2))
    assert torch.equal(TF.TensorArrayRead().forward(T(h2, torch.tensor(1), f2)), torch.tensor([2., 3., 4.]))

    stack = TF.StackCreator().forward(None)
    TF.StackPush().forward(T(stack, torch.tensor([1.])))
    TF.StackPush().forward(T(stack, torch.tensor([2.])))
    assert float(TF.StackPop().forward(T(stack))) == 2.0 and float(TF.StackPop().forward(T(stack))) == 1.0

    var = TF.Variable(torch.zeros(2, 2))
    ref = var.forward(torch.zeros(1))
    out = TF.Assign().forward(T(ref, torch.ones(2, 2)))
    assert torch.equal(out, torch.ones(2, 2)) and torch.equal(var.forward(torch.zeros(1)), torch.ones(2, 2))
    assert TF.ControlDependency().forward(T(torch.tensor([1.]))) is None
