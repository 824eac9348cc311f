"""pyspark/bigdl/examples/keras/imdb_cnn_lstm.py -> examples/keras_imdb_lstm.py of this repository."""
import os
import runpy
import sys

if __name__ == "__main__":
    here = os.path.dirname(os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
    sys.argv = [os.path.join(here, "examples", "keras_imdb_lstm.py")] + sys.argv[1:]
    runpy.run_path(sys.argv[0], run_name="__main__")
