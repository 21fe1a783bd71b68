"""ResNet stem convolution (4x4/1 over the 16-channel space-to-depth input) as a tcgen05
implicit GEMM (csrc/conv_stem.cu, K11): forward and weight gradient run the hand-written kernels
(the input is data, so there is no dgrad)."""
from __future__ import annotations

import torch
import torch.nn.functional as F

from . import _lib


USE_CUDNN_WGRAD = False      # test switch: weight gradient through cuDNN instead of the kernel


def _eligible(x: torch.Tensor, w: torch.Tensor) -> bool:
    if not (x.is_cuda and x.dtype == torch.bfloat16 and w.dtype == torch.bfloat16):
        return False
    if x.dim() != 4 or x.shape[1] != 16 or tuple(w.shape) != (64, 16, 4, 4):
        return False
    oh, ow = x.shape[2] - 3, x.shape[3] - 3
    return oh % 8 == 0 and ow % 16 == 0 and x.is_contiguous(memory_format=torch.channels_last)


class _StemConv(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w):
        n, _, ih, iw = x.shape
        oh, ow = ih - 3, iw - 3
        wk = w.contiguous(memory_format=torch.channels_last)       # [64, kh, kw, c] in memory
        y = torch.empty((n, 64, oh, ow), dtype=x.dtype, device=x.device,
                        memory_format=torch.channels_last)
        _lib.check(_lib.load().tfop_conv_stem_s2d_fwd(x.data_ptr(), wk.data_ptr(), y.data_ptr(),
                                                      n, oh, ow, _lib.stream_handle()),
                   "conv_stem_s2d_fwd")
        ctx.save_for_backward(x, w)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, w = ctx.saved_tensors
        dy = dy.contiguous(memory_format=torch.channels_last)
        n, _, oh, ow = dy.shape
        if USE_CUDNN_WGRAD:
            _, dw, _ = torch.ops.aten.convolution_backward(
                dy, x, w, None, [1, 1], [0, 0], [1, 1], False, [0, 0], 1, [False, True, False])
            return None, dw
        # fp32 [64, kh, kw, c], accumulated by the kernel's split-K RED -> logical [64,16,4,4]
        dw32 = torch.zeros((64, 4, 4, 16), dtype=torch.float32, device=dy.device)
        _lib.check(_lib.load().tfop_conv_stem_s2d_wgrad(x.data_ptr(), dy.data_ptr(), dw32.data_ptr(),
                                                        n, oh, ow, _lib.stream_handle()),
                   "conv_stem_s2d_wgrad")
        return None, dw32.permute(0, 3, 1, 2).to(w.dtype)


def stem_conv_s2d(x: torch.Tensor, w: torch.Tensor) -> torch.Tensor:
    """conv2d(x, w) for the space-to-depth stem: x logical [N,16,H,W] channels-last, w [64,16,4,4]."""
    if _eligible(x, w):
        return _StemConv.apply(x, w)
    if x.is_cuda:
        raise ValueError(
            f"stem_conv_s2d: the tcgen05 stem kernel takes bf16 channels-last [N,16,OH+3,OW+3] with "
            f"OH % 8 == 0 and OW % 16 == 0 (image sides that are multiples of 32), got "
            f"{tuple(x.shape)} {x.dtype}; feed the model the plain [N,3,H,W] input form instead")
    return F.conv2d(x, w)
