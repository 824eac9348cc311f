from bigdl_b200.interop.onnx_ops import (_sum, average_pool, batch_norm, concat, constant, conv, gather, gemm,  # noqa: F401
                                         max_pool, relu, reshape, shape, softmax, unsqueeze)
