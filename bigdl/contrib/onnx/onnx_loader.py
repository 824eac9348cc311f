from bigdl_b200.interop.onnx_loader import load, load_model, load_model_proto  # noqa: F401


class OnnxLoader:
    """pyspark/bigdl/contrib/onnx/onnx_loader.py ``OnnxLoader``: ``load_model(path)`` / ``load_graph(graph_proto)``."""

    def load_model(self, file_path):
        return load(file_path)

    def load_graph(self, graph_proto):
        from bigdl_b200.interop import onnx_pb as P
        m = P.ModelProto()
        m.graph.CopyFrom(graph_proto)
        return load_model_proto(m.SerializeToString())
