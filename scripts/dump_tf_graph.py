"""Print the nodes of a frozen TensorFlow GraphDef and flag the ops the loader does not support
(reference: scripts/dump_tf_graph.py, which needs a TensorFlow install; this one parses the protobuf directly)."""
import sys

sys.path.insert(0, __file__.rsplit("/", 2)[0])
from bigdl_b200.interop import tf_pb  # noqa: E402
from bigdl_b200.interop.tensorflow import SUPPORTED_OPS  # noqa: E402

gd = tf_pb.GraphDef()
with open(sys.argv[1], "rb") as f:
    gd.ParseFromString(f.read())
missing = set()
for n in gd.node:
    ok = n.op in SUPPORTED_OPS
    if not ok:
        missing.add(n.op)
    print(f"{'  ' if ok else '! '}{n.op:28s} {n.name}  <- {list(n.input)}")
print(f"\n{len(gd.node)} nodes; unsupported ops: {sorted(missing) or 'none'}")
