"""TensorFlow V2 checkpoints, .meta / SavedModel graphs and freezing, parsed without TensorFlow
(P/util/tf_utils.py, scripts/export_tf_checkpoint.py)."""
import os
import struct
import subprocess
import sys

import numpy as np
import pytest
import torch

from bigdl_b200.interop import tf_checkpoint as C
from bigdl_b200.interop import tf_pb as P
from bigdl_b200.interop.tensorflow import _GraphBuilder, load_tf

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _vars(n=300):
    rs = np.random.RandomState(0)
    v = {f"tower/layer_{i:03d}/kernel": rs.randn(3, 5).astype(np.float32) for i in range(n)}
    v.update({"global_step": np.array(7, np.int64), "stats/mean": rs.randn(4).astype(np.float64),
              "flags": np.array([True, False]), "ids": np.arange(6, dtype=np.int32).reshape(2, 3)})
    return v


def test_checkpoint_round_trip_multi_block(tmp_path):
    pre = str(tmp_path / "model.ckpt-7")
    v = _vars()
    assert C.write_checkpoint(pre, v) == pre
    assert os.path.getsize(pre + ".index") > 2 * 4096                       # several data blocks
    back = C.read_checkpoint(pre)
    assert set(back) == set(v)
    for k in v:
        assert back[k].dtype == v[k].dtype and back[k].shape == v[k].shape and np.array_equal(back[k], v[k]), k
    some = C.read_checkpoint(str(tmp_path), names=["global_step", "ids"])       # folder → the `checkpoint` state file
    assert set(some) == {"global_step", "ids"}
    listed = dict((n, (s, d)) for n, s, d in C.list_variables(pre + ".index"))
    assert listed["ids"] == ([2, 3], np.dtype(np.int32)) and listed["global_step"][0] == []
    assert C.resolve_prefix(pre + ".data-00000-of-00001") == pre and C.resolve_prefix(pre + ".meta") == pre
    # keys come back sorted and the footer carries the table magic
    keys = [k for k, _ in C.read_table(pre + ".index")]
    assert keys == sorted(keys) and keys[0] == b""
    raw = open(pre + ".index", "rb").read()
    assert struct.unpack("<Q", raw[-8:])[0] == 0xDB4775248B80FB57 and len(raw[-48:]) == 48


def test_corruption_is_detected(tmp_path):
    pre = str(tmp_path / "m")
    C.write_checkpoint(pre, {"w": np.arange(10, dtype=np.float32)})
    data = pre + ".data-00000-of-00001"
    b = bytearray(open(data, "rb").read())
    b[5] ^= 0xFF
    open(data, "wb").write(bytes(b))
    with pytest.raises(ValueError, match="tensor checksum mismatch"):
        C.read_checkpoint(pre)
    assert C.read_checkpoint(pre, verify=False)["w"].shape == (10,)
    idx = bytearray(open(pre + ".index", "rb").read())
    idx[3] ^= 0x55
    open(pre + ".index", "wb").write(bytes(idx))
    with pytest.raises(ValueError, match="block checksum mismatch"):
        C.read_checkpoint(pre)
    with pytest.raises(ValueError, match="bad magic"):
        C.read_table(data)


def test_snappy_blocks_and_masking():
    # literal "abcd" then a copy of 8 bytes at offset 4 (overlapping): "abcdabcdabcd"
    comp = bytes([12, (3 << 2) | 0]) + b"abcd" + bytes([((8 - 4) << 2) | 1, 4])
    assert C._snappy_decompress(comp) == b"abcdabcdabcd"
    for crc in (0, 1, 0xDEADBEEF, 0xFFFFFFFF):
        assert C._unmask(C._mask(crc)) == crc
    assert C._mask(0) == 0xA282EAD8


def _training_graph():
    """x → MatMul(w) → BiasAdd(b) → Relu, with the usual initialiser / saver clutter around the variables."""
    f32 = np.dtype(np.float32)
    b = _GraphBuilder()
    b.node("Placeholder", "x", [], dtype=f32)
    b.node("VariableV2", "dense/kernel", [], dtype=f32)
    b.node("VariableV2", "dense/bias", [], dtype=f32)
    b.node("Identity", "dense/kernel/read", ["dense/kernel"], T=f32)
    b.node("Identity", "dense/bias/read", ["dense/bias"], T=f32)
    b.const("init_w", np.zeros((4, 3), np.float32))
    b.node("Assign", "dense/kernel/Assign", ["dense/kernel", "init_w"], T=f32)
    b.node("NoOp", "init", ["^dense/kernel/Assign"])
    b.node("MatMul", "dense/MatMul", ["x", "dense/kernel/read"], T=f32, transpose_a=False, transpose_b=False)
    b.node("BiasAdd", "dense/BiasAdd", ["dense/MatMul", "dense/bias/read"], T=f32, data_format="NHWC")
    b.node("Relu", "out", ["dense/BiasAdd"], T=f32)
    b.node("VariableV2", "global_step", [], dtype=np.dtype(np.int64))
    return b.gd


def _ld(field, payload):
    return bytes([(field << 3) | 2]) + C._varint(len(payload)) + payload


def test_meta_graph_freeze_and_export_script(tmp_path):
    from bigdl.util import tf_utils

    gd = _training_graph()
    rs = np.random.RandomState(1)
    w, bias = rs.randn(4, 3).astype(np.float32), rs.randn(3).astype(np.float32)
    pre = str(tmp_path / "ck" / "model.ckpt")
    C.write_checkpoint(pre, {"dense/kernel": w, "dense/bias": bias, "global_step": np.array(3, np.int64)})
    # MetaGraphDef: meta_info_def (1) then graph_def (2)
    with open(pre + ".meta", "wb") as f:
        f.write(_ld(1, b"\x0a\x03tag") + _ld(2, gd.SerializeToString()))
    assert [n.name for n in C.read_graph_def(pre + ".meta").node] == [n.name for n in gd.node]
    deps = tf_utils.get_path("out", pre + ".meta")
    assert deps[-1] == "out" and "dense/kernel" in deps and "init" not in deps and deps.index("x") < deps.index("dense/MatMul")

    x = torch.randn(5, 4)
    want = torch.relu(x @ torch.from_numpy(w) + torch.from_numpy(bias))
    frozen = str(tmp_path / "frozen.pb")
    tf_utils.merge_checkpoint(pre + ".meta", pre, ["out"], frozen)
    nodes = {n.name: n.op for n in C.read_graph_def(frozen).node}
    assert nodes["dense/kernel"] == "Const" and "init" not in nodes and "global_step" not in nodes
    assert torch.allclose(load_tf(frozen, ["x"], ["out"]).forward(x), want, atol=1e-6)
    with pytest.raises(KeyError):
        C.freeze(gd, {"dense/kernel": w}, ["out"])

    out_dir = str(tmp_path / "exported")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "scripts", "export_tf_checkpoint.py"), pre, out_dir],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    assert "3 variables" in r.stdout and sorted(os.listdir(out_dir)) == ["model.bin", "model.pb"]
    m = tf_utils.convert(["x"], ["out"], pb_path=os.path.join(out_dir, "model.pb"), bin_file=os.path.join(out_dir, "model.bin"))
    assert torch.allclose(m.forward(x), want, atol=1e-6)
    assert set(tf_utils.export_checkpoint(str(tmp_path / "ck"))) == {"dense/kernel", "dense/bias", "global_step"}
    tf_utils.save_variable_bigdl({"dense/kernel:0": w}, str(tmp_path / "v.bin"))
    assert np.array_equal(np.load(str(tmp_path / "v.bin"))["dense/kernel"], w)

    # SavedModel folder: saved_model.pb { schema (1), meta_graphs (2) } + variables/variables.*
    sm = tmp_path / "saved"
    (sm / "variables").mkdir(parents=True)
    C.write_checkpoint(str(sm / "variables" / "variables"), {"dense/kernel": w, "dense/bias": bias})
    (sm / "saved_model.pb").write_bytes(b"\x08\x01" + _ld(2, _ld(2, gd.SerializeToString())))
    assert C.resolve_prefix(str(sm)).endswith(os.path.join("variables", "variables"))
    pb, binf = tf_utils.dump_model(str(tmp_path / "from_saved"), str(sm), ckpt_file=str(sm))
    m2 = load_tf(pb, ["x"], ["out"], binFile=binf)
    assert torch.allclose(m2.forward(x), want, atol=1e-6)


def test_resource_variables_become_identity():
    f32 = np.dtype(np.float32)
    b = _GraphBuilder()
    b.node("Placeholder", "x", [], dtype=f32)
    b.node("VarHandleOp", "w", [], dtype=f32)
    b.node("ReadVariableOp", "w/Read", ["w"], dtype=f32)
    b.node("Mul", "y", ["x", "w/Read"], T=f32)
    out = C.freeze(b.gd, {"w": np.array([2.0, 3.0], np.float32)}, ["y"])
    ops = {n.name: n.op for n in out.node}
    assert ops == {"x": "Placeholder", "w": "Const", "w/Read": "Identity", "y": "Mul"}
    g = load_tf(out.SerializeToString(), ["x"], ["y"])
    assert g.forward(torch.tensor([1.0, 1.0])).tolist() == [2.0, 3.0]
