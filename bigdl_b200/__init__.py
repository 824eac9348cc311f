"""bigdl_b200 — a Blackwell-native deep-learning framework with BigDL's capabilities.

PyTorch is the tensor/autograd/allocator substrate; the hot path runs on hand-written sm_100a kernels
(``csrc/*.cu`` → ``bigdl_b200/lib/libbigdl_b200.so``) and multi-GPU training runs one process per GPU with
parameters synchronised through NVLink peer memory (``bigdl_b200.parallel``).
"""
__version__ = "0.1.0"
