from __future__ import annotations

import os

import torch
import torch.distributed as dist


def build_model(name: str, class_num: int = 1000):
    name = name.lower()
    if name == "inception_v1":
        from ..models.inception import Inception_v1_NoAuxClassifier

        return Inception_v1_NoAuxClassifier(class_num), (3, 224, 224)
    if name == "inception_v2":
        from ..models.inception_v2 import Inception_v2_NoAuxClassifier

        return Inception_v2_NoAuxClassifier(class_num), (3, 224, 224)
    if name == "vgg16":
        from ..models.vgg import Vgg_16

        return Vgg_16(class_num), (3, 224, 224)
    if name == "vgg19":
        from ..models.vgg import Vgg_19

        return Vgg_19(class_num), (3, 224, 224)
    if name == "resnet50":
        from ..models.resnet import ResNet, model_init

        m = ResNet(class_num, depth=50, shortcut_type="B", dataset="imagenet", with_log_softmax=True)
        model_init(m)
        return m, (3, 224, 224)
    if name == "lenet5":
        from ..models.lenet import LeNet5

        return LeNet5(10), (1, 28, 28)
    raise ValueError(f"unknown model {name}")


def init_distributed():
    """Join the torchrun process group when launched under it; returns (rank, world, device)."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    cuda = torch.cuda.is_available()
    if cuda:
        torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
    if world > 1 and not dist.is_initialized():
        dist.init_process_group("nccl" if cuda else "gloo")
    return rank, world, torch.device("cuda", torch.cuda.current_device()) if cuda else torch.device("cpu")


def synthetic_batch(batch: int, shape, kind: str, class_num: int = 1000):
    x = torch.full((batch, *shape), 0.01) if kind == "constant" else torch.rand(batch, *shape)
    y = torch.ones(batch)           # label 1 for every record, as in the reference harness
    return x, y
