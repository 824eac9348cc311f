from bigdl_b200.models.datasets import *  # noqa: F401,F403
from bigdl_b200.models.datasets import (Cifar10DataSet, ImageNet2012, ImageNet2012Val, ImageNetDataSet,  # noqa: F401
                                        SequencePreprocess, toAutoencoderBatch)
