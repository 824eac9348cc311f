"""P/util/tf_utils.py: TensorFlow graph helpers. ``convert`` turns a frozen graph into a framework model; the helpers
that need a live TensorFlow session raise a clear error (TensorFlow is not installed in this environment)."""
from bigdl_b200.interop.tensorflow import load_tf, save_tf  # noqa: F401


def convert(input_ops, output_ops, byte_order="little_endian", bigdl_type="float", pb_path=None):
    if pb_path is None:
        raise NotImplementedError("freezing a live TensorFlow session needs TensorFlow; pass pb_path to a frozen graph")
    return load_tf(pb_path, input_ops, output_ops, byte_order)


def dump_model(path, graph=None, sess=None, ckpt_file=None, bigdl_type="float"):
    raise NotImplementedError("dump_model needs a TensorFlow runtime, which is not installed here")


def export_checkpoint(checkpoint_path):
    raise NotImplementedError("export_checkpoint needs a TensorFlow runtime, which is not installed here")


def save_variable_bigdl(tensors, target_path, bigdl_type="float"):
    import numpy as np

    np.savez(target_path, **{k: np.asarray(v) for k, v in tensors.items()})


def get_path(output_name, sess=None):
    """Names of the graph nodes an output depends on (the reference walks a live session's graph; pass a GraphDef)."""
    raise NotImplementedError("get_path needs a TensorFlow runtime, which is not installed here")


def merge_checkpoint(input_graph, checkpoint, output_node_names, output_graph, sess=None):
    """Freeze a checkpoint into a GraphDef (TensorFlow's freeze_graph); needs TensorFlow."""
    raise NotImplementedError("merge_checkpoint needs a TensorFlow runtime, which is not installed here")
