"""uint8 NHWC image batch -> normalised bf16 NHWC on the device (csrc/preproc.cu).

Per-step host feed of the reference trainers: examples/v1/dist-mnist/dist_mnist.py:275-281.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib

IMAGENET_MEAN = (0.485, 0.456, 0.406)
IMAGENET_STD = (0.229, 0.224, 0.225)


def u8_to_bf16_normalize(src_u8: torch.Tensor, out_bf16: torch.Tensor | None = None,
                         mean=IMAGENET_MEAN, std=IMAGENET_STD, stream=None) -> torch.Tensor:
    """src_u8: [N, H, W, C] uint8 (C<=4). Returns a logical [N, C, H, W] channels-last bf16."""
    n, h, w, c = src_u8.shape
    if out_bf16 is None:
        out_bf16 = torch.empty((n, h, w, c), dtype=torch.bfloat16, device=src_u8.device)
    if not src_u8.is_cuda:
        m = torch.tensor(mean, dtype=torch.float32)
        s = torch.tensor(std, dtype=torch.float32)
        out_bf16.copy_(((src_u8.float() / 255.0 - m) / s).to(torch.bfloat16))
        return out_bf16.permute(0, 3, 1, 2)
    lib = _lib.load()
    mean_a = (C.c_float * c)(*mean[:c])
    std_a = (C.c_float * c)(*std[:c])
    _lib.check(lib.tfop_u8_to_bf16_normalize(
        src_u8.data_ptr(), out_bf16.data_ptr(), src_u8.numel(), c, mean_a, std_a,
        _lib.stream_handle(stream)), "u8_to_bf16_normalize")
    return out_bf16.view(n, h, w, c).permute(0, 3, 1, 2)
