"""NHWC bf16 convolutions (1x1 / 3x3, stride 1 / 2) on the hand-written tcgen05 implicit-GEMM
kernels of csrc/conv_tc.cu: forward, data gradient and weight gradient.

Three autograd functions cover the ResNet blocks:

* ``conv2d(x, w, stride, stats)``            - one convolution;
* ``conv2d_skip(x, w, stride, stats)``       - returns ``(conv(x), x)``: the block input also feeds
  the identity branch, and in backward the convolution's data gradient is ADDED into the
  identity branch's gradient by the kernel's epilogue (TMA reduce-add) instead of by a separate
  elementwise kernel;
* ``conv2d_pair(x, w1, s1, st1, w2, s2, st2)`` - two convolutions of one input (the first conv of
  a block and its projection shortcut); the second data gradient is reduce-added into the first.

``stats``: a zeroed ``[2, Cout]`` fp32 tensor to which the kernel adds the per-channel sum and sum
of squares of its output - FusedBatchNorm2d takes it in place of its own statistics pass.

CPU tensors run ``F.conv2d`` so the models work in the CPU test-suite. CUDA tensors that the
kernels cannot take (not bf16, channels not a multiple of 64) raise: there is no silent library
fallback on the GPU.

Reference parity: the Conv2D layers of
examples/v1/distribution_strategy/keras-API/multi_worker_strategy-with-keras.py:42-60.
"""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import _lib

CL = torch.channels_last

# how often a shortcut gradient was accumulated inside the dgrad kernel (tests read this)
INPLACE_SKIP_ADDS = 0


def _check(x: torch.Tensor, w: torch.Tensor, stride: int) -> None:
    cout, cin, kh, kw = w.shape
    if x.dtype != torch.bfloat16 or w.dtype != torch.bfloat16:
        raise TypeError(f"tcgen05 convolution needs bf16 tensors, got {x.dtype} / {w.dtype}")
    if x.dim() != 4 or x.shape[1] != cin:
        raise ValueError(f"input {tuple(x.shape)} does not match weight {tuple(w.shape)}")
    if kh != kw or kh not in (1, 3) or stride not in (1, 2):
        raise ValueError(f"tcgen05 convolution covers 1x1 / 3x3, stride 1 / 2; got {kh}x{kw}/{stride}")
    if cin % 64 or cout % 64:
        raise ValueError(f"channel counts must be multiples of 64, got {cin} -> {cout}")


def _nhwc(t: torch.Tensor) -> torch.Tensor:
    return t if t.is_contiguous(memory_format=CL) else t.contiguous(memory_format=CL)


def _out_hw(h: int, w: int, k: int, stride: int):
    pad = k // 2
    return (h + 2 * pad - k) // stride + 1, (w + 2 * pad - k) // stride + 1


def conv_fwd(x, w, stride, stats=None):
    """y = conv2d(x, w, stride, padding=k//2) on the tcgen05 kernel (no autograd)."""
    _check(x, w, stride)
    x, w = _nhwc(x), _nhwc(w)
    n, cin, h, wd = x.shape
    cout, k = w.shape[0], w.shape[2]
    oh, ow = _out_hw(h, wd, k, stride)
    y = torch.empty((n, cout, oh, ow), dtype=x.dtype, device=x.device, memory_format=CL)
    _lib.check(_lib.load().tfop_conv_fwd(x.data_ptr(), w.data_ptr(), y.data_ptr(), _lib.ptr(stats),
                                         n, h, wd, cin, cout, k, stride, _lib.stream_handle()),
               "conv_fwd")
    return y


def conv_dgrad(dy, w, x_shape, stride, out=None, accumulate=False):
    """dx of conv_fwd; ``out`` + ``accumulate`` adds into an existing NHWC tensor."""
    n, cin, h, wd = x_shape
    cout, k = w.shape[0], w.shape[2]
    dy, w = _nhwc(dy), _nhwc(w)
    if out is None:
        out = torch.empty(tuple(x_shape), dtype=dy.dtype, device=dy.device, memory_format=CL)
        accumulate = False
    _lib.check(_lib.load().tfop_conv_dgrad(dy.data_ptr(), w.data_ptr(), out.data_ptr(), n, h, wd,
                                           cin, cout, k, stride, int(accumulate),
                                           _lib.stream_handle()), "conv_dgrad")
    return out


def conv_wgrad(x, dy, w, stride):
    """dw of conv_fwd, logical [Cout, Cin, k, k] in the weight's dtype and memory layout. A flat
    gradient bucket (parallel/flat.GradTarget) receives it directly when the engine set one."""
    n, cin, h, wd = x.shape
    cout, k = w.shape[0], w.shape[2]
    x, dy = _nhwc(x), _nhwc(dy)
    dw32 = torch.zeros((cout, k, k, cin), dtype=torch.float32, device=x.device)
    _lib.check(_lib.load().tfop_conv_wgrad(x.data_ptr(), dy.data_ptr(), dw32.data_ptr(), n, h, wd,
                                           cin, cout, k, stride, _lib.stream_handle()),
               "conv_wgrad")
    dw32 = dw32.permute(0, 3, 1, 2)
    target = getattr(w, "_tfop_grad_target", None)
    out = target.take() if target is not None else None
    if out is not None and out.dtype == w.dtype:
        out.copy_(dw32)
        return out
    return dw32.to(w.dtype)


def _usable_inplace(g, x_shape) -> bool:
    return (g is not None and g.dtype == torch.bfloat16 and tuple(g.shape) == tuple(x_shape)
            and g.is_contiguous(memory_format=CL) and g._base is None)


class _ConvFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, stride, stats):
        y = conv_fwd(x, w, stride, stats)
        ctx.save_for_backward(x, w)
        ctx.stride = stride
        return y

    @staticmethod
    def backward(ctx, dy):
        x, w = ctx.saved_tensors
        dx = conv_dgrad(dy, w, x.shape, ctx.stride) if ctx.needs_input_grad[0] else None
        dw = conv_wgrad(x, dy, w, ctx.stride) if ctx.needs_input_grad[1] else None
        return dx, dw, None, None


class _ConvSkipFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, stride, stats, inplace):
        y = conv_fwd(x, w, stride, stats)
        ctx.save_for_backward(x, w)
        ctx.stride, ctx.inplace = stride, inplace
        ctx.set_materialize_grads(False)
        return y, x.view_as(x)

    @staticmethod
    def backward(ctx, dy, dskip):
        x, w = ctx.saved_tensors
        dx = None
        if ctx.needs_input_grad[0]:
            if dy is None:
                dx = dskip
            elif ctx.inplace and _usable_inplace(dskip, x.shape):
                # the identity branch's gradient tensor becomes dx: dgrad is added into it
                global INPLACE_SKIP_ADDS
                INPLACE_SKIP_ADDS += 1
                dx = conv_dgrad(dy, w, x.shape, ctx.stride, out=dskip, accumulate=True)
            else:
                dx = conv_dgrad(dy, w, x.shape, ctx.stride)
                if dskip is not None:
                    dx = dx + dskip
        dw = conv_wgrad(x, dy, w, ctx.stride) if (ctx.needs_input_grad[1] and dy is not None) else None
        return dx, dw, None, None, None


class _ConvPairFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w1, s1, st1, w2, s2, st2):
        y1 = conv_fwd(x, w1, s1, st1)
        y2 = conv_fwd(x, w2, s2, st2)
        ctx.save_for_backward(x, w1, w2)
        ctx.s1, ctx.s2 = s1, s2
        ctx.set_materialize_grads(False)
        return y1, y2

    @staticmethod
    def backward(ctx, dy1, dy2):
        x, w1, w2 = ctx.saved_tensors
        dx = None
        if ctx.needs_input_grad[0]:
            if dy1 is not None:
                dx = conv_dgrad(dy1, w1, x.shape, ctx.s1)
            if dy2 is not None:
                dx = conv_dgrad(dy2, w2, x.shape, ctx.s2, out=dx, accumulate=dx is not None)
        dw1 = conv_wgrad(x, dy1, w1, ctx.s1) if (ctx.needs_input_grad[1] and dy1 is not None) else None
        dw2 = conv_wgrad(x, dy2, w2, ctx.s2) if (ctx.needs_input_grad[4] and dy2 is not None) else None
        return dx, dw1, None, None, dw2, None, None


def _native(x: torch.Tensor) -> bool:
    return x.is_cuda


def conv2d(x, w, stride=1, stats=None):
    if _native(x):
        return _ConvFn.apply(x, w, stride, stats)
    return F.conv2d(x, w, stride=stride, padding=w.shape[2] // 2)


def conv2d_skip(x, w, stride=1, stats=None, inplace_skip_grad=True):
    """``(conv(x), x)``; set ``inplace_skip_grad=False`` unless the consumer of the second output
    hands back a gradient tensor nobody else reads (FusedBatchNorm2d with ReLU does)."""
    if _native(x):
        return _ConvSkipFn.apply(x, w, stride, stats, inplace_skip_grad)
    return F.conv2d(x, w, stride=stride, padding=w.shape[2] // 2), x


def conv2d_pair(x, w1, s1, st1, w2, s2, st2):
    if _native(x):
        return _ConvPairFn.apply(x, w1, s1, st1, w2, s2, st2)
    return (F.conv2d(x, w1, stride=s1, padding=w1.shape[2] // 2),
            F.conv2d(x, w2, stride=s2, padding=w2.shape[2] // 2))


class TcConv2d(nn.Module):
    """Bias-free k x k convolution, padding k // 2, whose CUDA path is the tcgen05 kernel.
    Parameter name and shape match ``nn.Conv2d`` (``weight`` [Cout, Cin, k, k]), so state dicts
    interchange with the stock-module twin of the model."""

    def __init__(self, cin: int, cout: int, k: int, stride: int = 1):
        super().__init__()
        self.in_channels, self.out_channels, self.kernel_size, self.stride = cin, cout, k, stride
        self.weight = nn.Parameter(torch.empty(cout, cin, k, k))
        nn.init.kaiming_normal_(self.weight, mode="fan_out", nonlinearity="relu")
        # [2, Cout] sums the kernel's epilogue adds to and bn_finalize clears
        self.register_buffer("stat_sums", torch.zeros(2, cout, dtype=torch.float32),
                             persistent=False)

    def sums(self, x: torch.Tensor):
        """The statistics accumulator to pass as ``stats`` (None off-GPU)."""
        if not x.is_cuda:
            return None
        if self.stat_sums.dtype != torch.float32:      # a blanket .to(bfloat16) casts buffers too
            self.stat_sums = torch.zeros(2, self.out_channels, dtype=torch.float32, device=x.device)
        return self.stat_sums

    def forward(self, x, stats=None):
        return conv2d(x, self.weight, self.stride, stats)

    def extra_repr(self):
        return (f"{self.in_channels}, {self.out_channels}, kernel_size={self.kernel_size}, "
                f"stride={self.stride}")
