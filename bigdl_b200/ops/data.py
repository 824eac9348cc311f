"""GPU input-pipeline ops (csrc/data_ops.cu): the uint8 batch is uploaded raw (4x fewer PCIe bytes than fp32) and
cropped / flipped / normalised on the device in one pass. CPU tensors take the equivalent torch path, which is also the
oracle of the kernel test. Reference: S/dataset/image/{BGRImgCropper,HFlip,BGRImgNormalizer,BGRImgToBatch}.scala."""
from __future__ import annotations

import ctypes

import torch

from . import _lib


def image_batch(frames: torch.Tensor, crop_h: int, crop_w: int, mean, std, crop_xy: torch.Tensor | None = None,
                flip: torch.Tensor | None = None, dtype: torch.dtype = torch.bfloat16,
                channels_last: bool = True) -> torch.Tensor:
    """frames: uint8 [N, H, W, C] (C <= 4). crop_xy: int32 [N, 2] = (x0, y0) per image (None: centre crop); flip: uint8
    [N] (None: no flip). Returns a logical [N, C, crop_h, crop_w] tensor — channels_last memory (what the stem kernels
    read directly) or contiguous NCHW."""
    assert frames.dtype == torch.uint8 and frames.dim() == 4
    N, H, W, C = frames.shape
    if not frames.is_cuda:
        return _image_batch_torch(frames, crop_h, crop_w, mean, std, crop_xy, flip, dtype, channels_last)
    frames = frames.contiguous()
    lib = _lib.cuda_lib()
    m = (ctypes.c_float * C)(*[float(v) for v in mean])
    s = (ctypes.c_float * C)(*[float(v) for v in std])
    if channels_last:
        out = torch.empty((N, crop_h, crop_w, C), dtype=dtype, device=frames.device)
    else:
        out = torch.empty((N, C, crop_h, crop_w), dtype=dtype, device=frames.device)
    if crop_xy is not None:
        crop_xy = crop_xy.to(device=frames.device, dtype=torch.int32).contiguous()
    if flip is not None:
        flip = flip.to(device=frames.device, dtype=torch.uint8).contiguous()
    assert dtype in (torch.bfloat16, torch.float32)
    _lib.check(lib.bdl_image_batch_u8(frames.data_ptr(), out.data_ptr(), _lib.ptr(crop_xy), _lib.ptr(flip), N, H, W, C,
                                      crop_h, crop_w, m, s, int(dtype == torch.float32), int(not channels_last),
                                      _lib.stream_ptr()), "bdl_image_batch_u8")
    _lib.count_launch()
    return out.permute(0, 3, 1, 2) if channels_last else out


def _image_batch_torch(frames, crop_h, crop_w, mean, std, crop_xy, flip, dtype, channels_last):
    N, H, W, C = frames.shape
    outs = []
    mean_t = torch.tensor([float(v) for v in mean], dtype=torch.float32, device=frames.device)
    std_t = torch.tensor([float(v) for v in std], dtype=torch.float32, device=frames.device)
    for n in range(N):
        x0, y0 = ((W - crop_w) // 2, (H - crop_h) // 2) if crop_xy is None else (int(crop_xy[n, 0]), int(crop_xy[n, 1]))
        img = frames[n, y0:y0 + crop_h, x0:x0 + crop_w].float()
        if flip is not None and int(flip[n]):
            img = img.flip(1)
        outs.append((img - mean_t) / std_t)
    out = torch.stack(outs).to(dtype)                       # [N, h, w, C]
    return out.permute(0, 3, 1, 2) if channels_last else out.permute(0, 3, 1, 2).contiguous()


class GPUImageBatcher:
    """Turns decoded uint8 frames into a normalised training batch on the device: random (train) or centre (eval) crop,
    random horizontal flip, channel normalisation. The random draws come from the framework RNG so runs reproduce."""

    def __init__(self, crop_h: int, crop_w: int, mean, std, train: bool = True, dtype=torch.bfloat16):
        self.crop_h, self.crop_w, self.mean, self.std, self.train, self.dtype = crop_h, crop_w, mean, std, train, dtype

    def __call__(self, frames: torch.Tensor, device=None) -> torch.Tensor:
        from ..utils.random_generator import RNG

        N, H, W, _ = frames.shape
        crop = flip = None
        if self.train:
            g = torch.Generator().manual_seed(int(RNG.random()) & 0x7FFFFFFF)
            xs = torch.randint(0, W - self.crop_w + 1, (N,), generator=g, dtype=torch.int32)
            ys = torch.randint(0, H - self.crop_h + 1, (N,), generator=g, dtype=torch.int32)
            crop = torch.stack([xs, ys], dim=1)
            flip = (torch.rand(N, generator=g) < 0.5).to(torch.uint8)
        if device is not None and frames.device != torch.device(device):
            frames = frames.to(device, non_blocking=True)
        return image_batch(frames, self.crop_h, self.crop_w, self.mean, self.std, crop, flip, self.dtype)
