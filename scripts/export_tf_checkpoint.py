"""Export the graph and the variables of a TensorFlow checkpoint (or a SavedModel folder) as ``model.pb`` +
``model.bin``, the pair ``Module.loadTF(pb, inputs, outputs, binFile=bin)`` reads — the job of the reference's
scripts/export_tf_checkpoint.py, done by parsing the checkpoint files directly (no TensorFlow install needed):

    python scripts/export_tf_checkpoint.py checkpoint_name            # reads checkpoint_name.meta / .index / .data-*
    python scripts/export_tf_checkpoint.py saver_folder
    python scripts/export_tf_checkpoint.py checkpoint_name save_path
    python scripts/export_tf_checkpoint.py saver_folder save_path
    python scripts/export_tf_checkpoint.py meta_file checkpoint_name save_path

The default save_path is ``model``."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main(argv=None):
    argv = list(sys.argv[1:] if argv is None else argv)
    from bigdl_b200.interop import tf_checkpoint as C

    if len(argv) in (1, 2):
        save_path = argv[1] if len(argv) == 2 else "model"
        if os.path.isdir(argv[0]):
            graph = argv[0] if os.path.exists(os.path.join(argv[0], "saved_model.pb")) else None
            ckpt = C.resolve_prefix(argv[0])
            if graph is None:
                graph = ckpt + ".meta"
        else:
            graph, ckpt = argv[0] + ".meta", argv[0]
    elif len(argv) == 3:
        graph, ckpt, save_path = argv
    else:
        print(__doc__)
        return 1
    if os.path.isfile(save_path):
        print("The save folder is a file. Exit")
        return 1
    pb, binf = C.dump_model(save_path, graph, ckpt)
    names = C.list_variables(ckpt)
    print(f"wrote {pb} and {binf} ({len(names)} variables)")
    return 0


if __name__ == "__main__":
    sys.exit(main())
