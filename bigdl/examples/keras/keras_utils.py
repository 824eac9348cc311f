"""pyspark/bigdl/examples/keras/keras_utils.py: save / load a Keras definition (+ weights) for the converter."""
import json


def save_keras_definition(keras_model, path):
    """Write ``keras_model.to_json()`` to ``path`` (accepts anything with ``to_json``)."""
    with open(path, "w") as f:
        f.write(keras_model.to_json() if hasattr(keras_model, "to_json") else json.dumps(keras_model))


def dump_keras(keras_model, dump_weights=False, path_prefix="/tmp/keras_model"):
    """-> (definition json path, hdf5 weight path or None), the pair ``Model.load_keras`` takes."""
    def_path = path_prefix + ".json"
    save_keras_definition(keras_model, def_path)
    weights_path = None
    if dump_weights:
        weights_path = path_prefix + ".h5"
        keras_model.save_weights(weights_path, overwrite=True)
    return def_path, weights_path
