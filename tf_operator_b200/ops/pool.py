"""MaxPool 3x3 / stride 2 / pad 1 for channels-last bf16 (csrc/pool.cu); torch elsewhere.

No counterpart in the reference (it ships no kernels).
"""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import _lib

class _MaxPool3x3s2(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        lib = _lib.load()
        if not x.permute(0, 2, 3, 1).is_contiguous():
            x = x.contiguous(memory_format=torch.channels_last)
        n, c, h, w = x.shape
        ho, wo = (h + 1) // 2, (w + 1) // 2
        y = torch.empty((n, c, ho, wo), dtype=x.dtype, device=x.device,
                        memory_format=torch.channels_last)
        idx = torch.empty(n * ho * wo * c, dtype=torch.uint8, device=x.device)
        _lib.check(lib.tfop_maxpool3x3s2_fwd(x.data_ptr(), y.data_ptr(), idx.data_ptr(),
                                             n, h, w, c, _lib.stream_handle()), "maxpool_fwd")
        ctx.save_for_backward(idx)
        ctx.xshape = (n, c, h, w)
        return y

    @staticmethod
    def backward(ctx, dy):
        lib = _lib.load()
        (idx,) = ctx.saved_tensors
        n, c, h, w = ctx.xshape
        if not dy.permute(0, 2, 3, 1).is_contiguous():
            dy = dy.contiguous(memory_format=torch.channels_last)
        dx = torch.empty((n, c, h, w), dtype=dy.dtype, device=dy.device,
                         memory_format=torch.channels_last)
        _lib.check(lib.tfop_maxpool3x3s2_bwd(dy.data_ptr(), idx.data_ptr(), dx.data_ptr(),
                                             n, h, w, c, _lib.stream_handle()), "maxpool_bwd")
        return dx


def max_pool_3x3_s2(x: torch.Tensor) -> torch.Tensor:
    if x.is_cuda and x.dtype == torch.bfloat16 and x.dim() == 4 and x.shape[1] % 8 == 0:
        return _MaxPool3x3s2.apply(x)
    return F.max_pool2d(x, 3, 2, 1)


class MaxPool3x3s2(nn.Module):
    def forward(self, x):
        return max_pool_3x3_s2(x)
