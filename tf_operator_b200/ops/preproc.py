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


def stem_space_to_depth(src_u8: torch.Tensor, out_bf16: torch.Tensor | None = None,
                        mean=IMAGENET_MEAN, std=IMAGENET_STD, stream=None) -> torch.Tensor:
    """uint8 [N,H,W,3] -> logical [N,16,H/2+3,W/2+3] channels-last bf16 (see csrc/preproc.cu).
    ``stem_weight_s2d`` gives the matching 4x4 weight of a 7x7/2/pad-3 convolution."""
    n, h, w, ch = src_u8.shape
    assert ch == 3 and h % 2 == 0 and w % 2 == 0
    ho, wo = h // 2 + 3, w // 2 + 3
    if out_bf16 is None:
        out_bf16 = torch.empty((n, ho, wo, 16), dtype=torch.bfloat16, device=src_u8.device)
    if not src_u8.is_cuda:
        m = torch.tensor(mean, dtype=torch.float32)
        s = torch.tensor(std, dtype=torch.float32)
        x = (src_u8.float() / 255.0 - m) / s                                  # [N,H,W,3]
        x = x.view(n, h // 2, 2, w // 2, 2, 3).permute(0, 1, 3, 2, 4, 5).reshape(n, h // 2, w // 2, 12)
        out_bf16.zero_()
        out_bf16[:, 2:2 + h // 2, 2:2 + w // 2, :12] = x.to(torch.bfloat16)
        return out_bf16.permute(0, 3, 1, 2)
    lib = _lib.load()
    mean_a = (C.c_float * 3)(*mean[:3])
    std_a = (C.c_float * 3)(*std[:3])
    _lib.check(lib.tfop_stem_space_to_depth(
        src_u8.data_ptr(), out_bf16.data_ptr(), n, h, w, mean_a, std_a,
        _lib.stream_handle(stream)), "stem_space_to_depth")
    return out_bf16.view(n, ho, wo, 16).permute(0, 3, 1, 2)


def stem_weight_s2d(w7: torch.Tensor) -> torch.Tensor:
    """[Cout,3,7,7] (stride 2, pad 3) -> [Cout,16,4,4] (stride 1, pad 0) for the s2d input.
    Differentiable, so the 7x7 parameter keeps receiving its gradient."""
    cout = w7.shape[0]
    w8 = torch.nn.functional.pad(w7, (1, 0, 1, 0))                 # zero tap in front: kh' = kh+1
    w8 = w8.reshape(cout, 3, 4, 2, 4, 2)                           # [o, ch, ah, bh, aw, bw]
    w4 = w8.permute(0, 3, 5, 1, 2, 4).reshape(cout, 12, 4, 4)      # channel = (bh*2+bw)*3+ch
    w4 = torch.nn.functional.pad(w4, (0, 0, 0, 0, 0, 4))           # 12 -> 16 channels
    return w4.contiguous(memory_format=torch.channels_last)
