"""The reference's ONNX extension point (P/contrib/onnx/{ops_converter,ops_mapping,converter_utils}.py): node-level
converters ``fn(inputs, prev_modules, attrs, outputs) -> (node, [shapes])`` reachable through the ``bigdl`` package."""
import numpy as np
import torch

from bigdl.contrib.onnx.converter_utils import calc_output_shape
from bigdl.contrib.onnx.ops_mapping import _convert_map
from bigdl.contrib.onnx import ops_converter as oc


def test_calc_output_shape():
    assert calc_output_shape(224, 7, padding=3, stride=2) == 112
    assert calc_output_shape(112, 3, padding=0, stride=2, ceil_mode=True) == 56
    assert calc_output_shape(10, 3, dilation=2) == 6


def test_convert_map_builds_a_runnable_graph():
    from bigdl_b200.nn.layer import Input
    from bigdl_b200.nn.module import Graph
    assert set(_convert_map) >= {"Conv", "BatchNormalization", "Relu", "MaxPool", "AveragePool", "Sum", "Reshape", "Gemm",
                                 "Softmax", "Shape", "Gather", "Unsqueeze", "Concat", "Constant"}
    rs = np.random.RandomState(0)
    w = rs.randn(4, 3, 3, 3).astype(np.float32)
    b = rs.randn(4).astype(np.float32)
    fc = rs.randn(4 * 4 * 4, 5).astype(np.float32)
    x_in = Input()
    shape = (2, 3, 8, 8)
    node, (shape,) = _convert_map["Conv"]([(None, shape), (w, w.shape), (b, b.shape)], [x_in],
                                          {"kernel_shape": (3, 3), "pads": (1, 1, 1, 1), "strides": (1, 1)}, None)
    assert shape == (2, 4, 8, 8)
    bn_in = [(None, shape)] + [(v, (4,)) for v in (np.ones(4, np.float32), np.zeros(4, np.float32),
                                                   np.zeros(4, np.float32), np.ones(4, np.float32))]
    node, (shape,) = oc.batch_norm(bn_in, [node], {"epsilon": 1e-5}, None)
    node, (shape,) = oc.relu([(None, shape)], [node], {}, None)
    node, (shape,) = oc.max_pool([(None, shape)], [node], {"kernel_shape": (2, 2), "strides": (2, 2)}, None)
    assert shape == (2, 4, 4, 4)
    node, _ = oc.reshape([(None, shape), (np.array([2, 64]), (2,))], [node], {}, None)
    node, (shape,) = oc.gemm([(None, (2, 64)), (fc, fc.shape)], [node], {}, None)
    assert shape == (2, 5)
    node, _ = oc.softmax([(None, shape)], [node], {}, None)
    model = Graph([x_in], [node]).evaluate()
    x = torch.from_numpy(rs.randn(2, 3, 8, 8).astype(np.float32))
    y = model.forward(x)
    ref = torch.nn.functional.conv2d(x, torch.from_numpy(w), torch.from_numpy(b), padding=1)
    ref = torch.relu(ref / (1 + 1e-5) ** 0.5)
    ref = torch.nn.functional.max_pool2d(ref, 2).reshape(2, 64) @ torch.from_numpy(fc)
    assert torch.allclose(y, torch.softmax(ref, -1), atol=1e-4)


def test_concat_sum_unsqueeze_shapes():
    from bigdl_b200.nn.layer import Input
    from bigdl_b200.nn.module import Graph
    a, b = Input(), Input()
    node, (shape,) = oc.concat([(None, (2, 3, 4)), (None, (2, 5, 4))], [a, b], {"axis": 1}, None)
    assert shape == (2, 8, 4)
    node2, (s2,) = oc.unsqueeze([(None, shape)], [node], {"axes": (0,)}, None)
    assert s2 == (1, 2, 8, 4)
    y = Graph([a, b], [node2]).forward([torch.ones(2, 3, 4), torch.zeros(2, 5, 4)])
    assert tuple(y.shape) == (1, 2, 8, 4) and float(y.sum()) == 24.0
    s, _ = oc._sum([(None, (2, 3)), (None, (2, 3))], [a, b], {}, None)
    z = Graph([a, b], [s]).forward([torch.ones(2, 3), torch.ones(2, 3)])
    assert float(z.sum()) == 12.0
