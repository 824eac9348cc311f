from bigdl_b200.models.inception import (Inception_Layer_v1 as inception_layer_v1, Inception_v1 as inception_v1,  # noqa: F401
                                         Inception_v1_NoAuxClassifier as inception_v1_no_aux_classifier)
