"""Convert a dict of checkpoint variables saved as .npz into the ``binFile`` accepted by ``Module.loadTF``
(reference: scripts/export_tf_checkpoint.py dumps a TF checkpoint; without TensorFlow the input here is an .npz whose
keys are the variable node names)."""
import sys

import numpy as np

src, dst = sys.argv[1], sys.argv[2]
data = dict(np.load(src))
np.savez(dst, **{k.split(":")[0]: np.asarray(v, dtype=np.float32) for k, v in data.items()})
print(f"wrote {len(data)} variables to {dst}")
