"""P/util/tf_utils.py: TensorFlow graph / checkpoint helpers. The reference drives a live TensorFlow session; here the
same results come from the files themselves (frozen ``.pb``, ``.meta``, SavedModel folders, V2 checkpoints) through
``bigdl_b200.interop.tf_checkpoint`` — no TensorFlow install is needed. Where the reference takes a ``sess`` /
``graph`` object, pass the path of the corresponding file instead."""
import os
import tempfile

import numpy as np

from bigdl_b200.interop import tf_checkpoint as _ck
from bigdl_b200.interop.tensorflow import load_tf, save_tf  # noqa: F401


def get_path(output_name, sess=None):
    """Names of the graph nodes ``output_name`` depends on. ``sess``: a GraphDef or the path of a ``.pb`` / ``.meta``."""
    gd = _ck.read_graph_def(sess) if isinstance(sess, str) else sess
    if gd is None:
        raise ValueError("get_path needs the graph: pass a GraphDef or the path of a .pb / .meta file as `sess`")
    return _ck.get_path(output_name, gd)


def convert(input_ops, output_ops, byte_order="little_endian", bigdl_type="float", pb_path=None, bin_file=None):
    """Frozen graph (or graph + exported variables) → model. ``input_ops`` / ``output_ops`` are node names."""
    if pb_path is None:
        raise ValueError("convert needs the graph file: pass pb_path (a frozen .pb, or model.pb with bin_file=model.bin)")
    return load_tf(pb_path, list(input_ops), list(output_ops), byte_order, bin_file)


def export_checkpoint(checkpoint_path):
    """{variable name: ndarray} of a checkpoint prefix / folder."""
    return _ck.read_checkpoint(checkpoint_path)


def save_variable_bigdl(tensors, target_path, bigdl_type="float"):
    """Write {name: array} in the ``binFile`` format of ``Module.loadTF``."""
    with open(target_path, "wb") as f:
        np.savez(f, **{k.split(":")[0]: np.asarray(v) for k, v in tensors.items()})


def dump_model(path, graph=None, sess=None, ckpt_file=None, bigdl_type="float"):
    """``path/model.pb`` + ``path/model.bin`` from a graph file (``graph``: .pb / .meta / SavedModel folder) and a
    checkpoint (``ckpt_file``; defaults to the checkpoint next to a ``.meta``)."""
    graph = graph if graph is not None else sess
    if ckpt_file is None and isinstance(graph, str) and graph.endswith(".meta"):
        ckpt_file = graph[:-5]
    return _ck.dump_model(path, graph, ckpt_file)


def merge_checkpoint(input_graph, checkpoint, output_node_names, output_graph, sess=None):
    """Freeze: variables → constants, keep only what ``output_node_names`` need, write ``output_graph``."""
    gd = _ck.read_graph_def(input_graph) if isinstance(input_graph, str) else input_graph
    frozen = _ck.freeze(gd, _ck.read_checkpoint(checkpoint), list(output_node_names))
    with open(output_graph, "wb") as f:
        f.write(frozen.SerializeToString())
    return output_graph
