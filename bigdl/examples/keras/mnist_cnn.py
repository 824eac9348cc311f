"""pyspark/bigdl/examples/keras/mnist_cnn.py -> examples/keras_mnist_cnn.py of this repository."""
import os
import runpy
import sys

if __name__ == "__main__":
    here = os.path.dirname(os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
    sys.argv = [os.path.join(here, "examples", "keras_mnist_cnn.py")] + sys.argv[1:]
    runpy.run_path(sys.argv[0], run_name="__main__")
