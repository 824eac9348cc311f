"""Model interop: Torch7 .t7 (T/utils/TorchFileSpec.scala style round trips)."""
import os
import struct

import numpy as np
import pytest
import torch

from bigdl_b200.interop import torchfile as T7
from bigdl_b200.nn import layer as L
from bigdl_b200.utils.table import Table


def test_t7_primitives_and_tensors(tmp_path):
    p = str(tmp_path / "a.t7")
    t = torch.arange(24, dtype=torch.float32).reshape(2, 3, 4)
    tab = Table()
    tab["x"], tab["flag"], tab["n"], tab["name"], tab[1] = t, True, 3.5, "hello", torch.tensor([1, 2, 3])
    T7.save(tab, p)
    back = T7.load(p)
    assert torch.equal(back["x"], t) and back["flag"] is True and back["n"] == 3.5 and back["name"] == "hello"
    assert back[1].dtype == torch.int64 and back[1].tolist() == [1, 2, 3]
    with pytest.raises(FileExistsError):
        T7.save(tab, p)
    # header layout: TYPE_TABLE, index 1, 5 entries
    assert struct.unpack("<iii", open(p, "rb").read(12)) == (3, 1, 5)
    # a non-contiguous tensor is stored by value
    T7.save(t.transpose(0, 2), p, overWrite=True)
    assert torch.equal(T7.load(p), t.transpose(0, 2))


def test_t7_module_roundtrip(tmp_path):
    torch.manual_seed(0)
    m = (L.Sequential()
         .add(L.SpatialConvolution(3, 8, 3, 3, 1, 1, 1, 1))
         .add(L.ReLU())
         .add(L.SpatialMaxPooling(2, 2, 2, 2))
         .add(L.SpatialCrossMapLRN(5, 1e-4, 0.75, 2.0))
         .add(L.Concat(2).add(L.Sequential().add(L.SpatialConvolution(8, 4, 1, 1)))
              .add(L.Sequential().add(L.SpatialConvolution(8, 4, 3, 3, 1, 1, 1, 1)).add(L.Threshold(0.1, 0.0))))
         .add(L.View(8 * 4 * 4))
         .add(L.Dropout(0.3))
         .add(L.Linear(128, 10))
         .add(L.LogSoftMax()))
    m.evaluate()
    x = torch.randn(2, 3, 8, 8)
    ref = m.forward(x).clone()
    p = str(tmp_path / "m.t7")
    m.saveTorch(p)
    m2 = L.Module.loadTorch(p).evaluate()
    assert [type(c).__name__ for c in m2.modules] == [type(c).__name__ for c in m.modules]
    assert torch.allclose(m2.forward(x), ref, atol=1e-6)
    with pytest.raises(ValueError):
        L.Sequential().add(L.SpatialConvolution(4, 4, 3, 3, n_group=2)).saveTorch(str(tmp_path / "g.t7"))


def test_t7_reads_torch_written_layout(tmp_path):
    """Hand-assemble the bytes Torch7 itself would write for nn.Linear (incl. a shared-storage back reference)."""
    out = bytearray()
    i32 = lambda v: out.extend(struct.pack("<i", v))
    i64 = lambda v: out.extend(struct.pack("<q", v))
    def s(v):
        i32(len(v)); out.extend(v.encode())
    def string(v):
        i32(2); s(v)
    i32(4); i32(1); s("V 1"); s("nn.Linear")
    i32(3); i32(2); i32(3)                       # table, index 2, 3 entries
    string("_type"); string("torch.FloatTensor")
    string("weight")
    i32(4); i32(3); s("V 1"); s("torch.FloatTensor")
    i32(2); i64(2); i64(3); i64(3); i64(1); i64(1)
    i32(4); i32(4); s("V 1"); s("torch.FloatStorage"); i64(8)
    out.extend(np.arange(8, dtype=np.float32).tobytes())
    string("bias")
    i32(4); i32(5); s("V 1"); s("torch.FloatTensor")
    i32(1); i64(2); i64(1); i64(7)               # bias = storage[6:8] of the SAME storage (index 4)
    i32(4); i32(4)
    p = tmp_path / "lin.t7"
    p.write_bytes(bytes(out))
    m = T7.load(str(p))
    assert isinstance(m, L.Linear) and m.weight.tolist() == [[0, 1, 2], [3, 4, 5]] and m.bias.tolist() == [6, 7]
