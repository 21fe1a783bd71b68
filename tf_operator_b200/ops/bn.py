"""Fused BatchNorm(+residual)(+ReLU) for channels-last bf16 activations.

CUDA tensors go through csrc/bn.cu (two HBM passes forward, two backward,
1 bit/element ReLU mask); CPU tensors use the plain PyTorch formula so models
still run in the CPU test-suite. No counterpart in the reference (it ships no
kernels).
"""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import _lib

_workspaces: dict = {}


def _workspace(device) -> torch.Tensor:
    key = (device.type, device.index)
    ws = _workspaces.get(key)
    if ws is None:
        nbytes = _lib.load().tfop_bn_workspace_bytes()
        ws = torch.zeros(nbytes, dtype=torch.uint8, device=device)
        _workspaces[key] = ws
    return ws


def _is_nhwc(x: torch.Tensor) -> bool:
    return x.dim() == 4 and x.permute(0, 2, 3, 1).is_contiguous()


def supported(x: torch.Tensor) -> bool:
    c = x.shape[1]
    return (x.is_cuda and x.dtype == torch.bfloat16 and x.dim() == 4 and c % 64 == 0
            and 64 <= c <= 4096)


class _FusedBNAct(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, res, gamma, beta, running_mean, running_var, training, momentum, eps,
                relu, sums=None):
        lib = _lib.load()
        if not _is_nhwc(x):
            x = x.contiguous(memory_format=torch.channels_last)
        if res is not None and not _is_nhwc(res):
            res = res.contiguous(memory_format=torch.channels_last)
        n, c, h, w = x.shape
        m = n * h * w
        st = _lib.stream_handle()
        pd = _lib.dtype_tag(gamma.dtype)
        dev = x.device
        scale_shift = torch.empty(2 * c, dtype=torch.float32, device=dev)
        if training:
            save_mean = torch.empty(c, dtype=torch.float32, device=dev)
            save_invstd = torch.empty(c, dtype=torch.float32, device=dev)
            if sums is not None:
                # the producing convolution's epilogue already reduced sum / sum of squares
                _lib.check(lib.tfop_bn_finalize(
                    sums.data_ptr(), m, c, gamma.data_ptr(), beta.data_ptr(), pd,
                    _lib.ptr(running_mean), _lib.ptr(running_var), float(momentum), float(eps),
                    scale_shift.data_ptr(), save_mean.data_ptr(), save_invstd.data_ptr(), st),
                    "bn_finalize")
            else:
                _lib.check(lib.tfop_bn_fwd_stats(
                    x.data_ptr(), m, c, gamma.data_ptr(), beta.data_ptr(), pd,
                    _lib.ptr(running_mean), _lib.ptr(running_var), float(momentum), float(eps),
                    scale_shift.data_ptr(), save_mean.data_ptr(), save_invstd.data_ptr(),
                    _workspace(dev).data_ptr(), st), "bn_fwd_stats")
        else:
            invstd = torch.rsqrt(running_var.float() + eps)
            a = gamma.float() * invstd
            scale_shift[:c] = a
            scale_shift[c:] = beta.float() - running_mean.float() * a
            save_mean, save_invstd = running_mean.float(), invstd
        y = torch.empty_like(x, memory_format=torch.channels_last)
        need_grad = training and (x.requires_grad or gamma.requires_grad)
        mask = (torch.empty(m * c // 8, dtype=torch.uint8, device=dev)
                if (relu and need_grad) else None)
        _lib.check(lib.tfop_bn_fwd_apply(
            x.data_ptr(), _lib.ptr(res), y.data_ptr(), _lib.ptr(mask), scale_shift.data_ptr(),
            m, c, int(relu), st), "bn_fwd_apply")
        ctx.save_for_backward(x, gamma, save_mean, save_invstd, mask)
        ctx.has_res = res is not None
        ctx.relu = relu
        ctx.training = training
        return y

    @staticmethod
    def backward(ctx, dy):
        lib = _lib.load()
        x, gamma, save_mean, save_invstd, mask = ctx.saved_tensors
        if not ctx.training:
            raise RuntimeError("FusedBatchNorm backward is only implemented for training mode")
        if not _is_nhwc(dy):
            dy = dy.contiguous(memory_format=torch.channels_last)
        n, c, h, w = x.shape
        m = n * h * w
        st = _lib.stream_handle()
        pd = _lib.dtype_tag(gamma.dtype)
        dev = x.device
        dgamma = torch.empty_like(gamma)
        dbeta = torch.empty_like(gamma)
        coef = torch.empty(2 * c, dtype=torch.float32, device=dev)
        _lib.check(lib.tfop_bn_bwd_reduce(
            dy.data_ptr(), x.data_ptr(), _lib.ptr(mask), save_mean.data_ptr(),
            save_invstd.data_ptr(), m, c, dgamma.data_ptr(), dbeta.data_ptr(), pd,
            coef.data_ptr(), _workspace(dev).data_ptr(), st), "bn_bwd_reduce")
        dx = torch.empty_like(x, memory_format=torch.channels_last)
        if ctx.has_res:
            # without ReLU the residual gradient is dy itself
            dres = torch.empty_like(x, memory_format=torch.channels_last) if ctx.relu else None
        else:
            dres = None
        _lib.check(lib.tfop_bn_bwd_apply(
            dy.data_ptr(), x.data_ptr(), _lib.ptr(mask), save_mean.data_ptr(),
            save_invstd.data_ptr(), gamma.data_ptr(), pd, coef.data_ptr(), dx.data_ptr(),
            _lib.ptr(dres), m, c, st), "bn_bwd_apply")
        if ctx.has_res and dres is None:
            dres = dy
        return dx, dres, dgamma, dbeta, None, None, None, None, None, None, None


def fused_bn_act(x, gamma, beta, running_mean, running_var, training, momentum=0.1, eps=1e-5,
                 relu=True, residual=None, sums=None):
    """y = [relu](batch_norm(x) [+ residual]); CUDA bf16 NHWC uses the fused kernels. ``sums``:
    the [2, C] fp32 sum / sum-of-squares array the producing convolution filled (ops/conv.py);
    it replaces the statistics pass over x and is cleared for the next step."""
    if supported(x):
        return _FusedBNAct.apply(x, residual, gamma, beta, running_mean, running_var, training,
                                 momentum, eps, relu, sums if training else None)
    # CPU tensors, or a channel count the kernels do not cover (not a multiple of 64)
    y = F.batch_norm(x, running_mean, running_var, gamma, beta, training, momentum, eps)
    if residual is not None:
        y = y + residual
    return F.relu(y) if relu else y


# backward of the fused stem tail: "gather" = both BN backward passes gather the pooled gradient on
# the fly (no dz tensor); "scatter" = one pool-backward kernel writes dz, then the plain BN passes
STEM_BWD_MODE = "scatter"    # measured: gather 0.76 + 0.78 ms, scatter 0.37 + 0.15 + 0.21 ms


class _BnReluMaxPool(torch.autograd.Function):
    """maxpool3x3/2/pad1(relu(bn(x))) for the ResNet stem without materialising the BN output
    (csrc/bn.cu: stem_bn_relu_pool_fwd, stem_bn_pool_bwd_{reduce,apply})."""

    @staticmethod
    def forward(ctx, x, gamma, beta, running_mean, running_var, momentum, eps):
        lib = _lib.load()
        if not _is_nhwc(x):
            x = x.contiguous(memory_format=torch.channels_last)
        n, c, h, w = x.shape
        ho, wo = (h + 1) // 2, (w + 1) // 2
        st = _lib.stream_handle()
        dev = x.device
        scale_shift = torch.empty(2 * c, dtype=torch.float32, device=dev)
        stats = torch.empty((2, c), dtype=torch.float32, device=dev)
        _lib.check(lib.tfop_bn_fwd_stats(
            x.data_ptr(), n * h * w, c, gamma.data_ptr(), beta.data_ptr(), _lib.dtype_tag(gamma.dtype),
            _lib.ptr(running_mean), _lib.ptr(running_var), float(momentum), float(eps),
            scale_shift.data_ptr(), stats[0].data_ptr(), stats[1].data_ptr(),
            _workspace(dev).data_ptr(), st), "bn_fwd_stats")
        y = torch.empty((n, c, ho, wo), dtype=x.dtype, device=dev, memory_format=torch.channels_last)
        idx = torch.empty(n * ho * wo * c, dtype=torch.uint8, device=dev)
        _lib.check(lib.tfop_stem_bn_relu_pool_fwd(x.data_ptr(), scale_shift.data_ptr(), y.data_ptr(),
                                                  idx.data_ptr(), n, h, w, c, st),
                   "stem_bn_relu_pool_fwd")
        ctx.save_for_backward(x, gamma, stats, idx)
        return y

    @staticmethod
    def backward(ctx, dy):
        lib = _lib.load()
        x, gamma, stats, idx = ctx.saved_tensors
        if not _is_nhwc(dy):
            dy = dy.contiguous(memory_format=torch.channels_last)
        n, c, h, w = x.shape
        st = _lib.stream_handle()
        pd = _lib.dtype_tag(gamma.dtype)
        dev = x.device
        dgamma, dbeta = torch.empty_like(gamma), torch.empty_like(gamma)
        coef = torch.empty(2 * c, dtype=torch.float32, device=dev)
        if STEM_BWD_MODE == "scatter":
            # materialise dz once with the pool gather kernel (dead code 15 matches no tap), then
            # the plain BN backward passes without a ReLU mask
            dz = torch.empty_like(x, memory_format=torch.channels_last)
            _lib.check(lib.tfop_maxpool3x3s2_bwd(dy.data_ptr(), idx.data_ptr(), dz.data_ptr(),
                                                 n, h, w, c, st), "maxpool_bwd")
            m = n * h * w
            _lib.check(lib.tfop_bn_bwd_reduce(
                dz.data_ptr(), x.data_ptr(), 0, stats[0].data_ptr(), stats[1].data_ptr(), m, c,
                dgamma.data_ptr(), dbeta.data_ptr(), pd, coef.data_ptr(),
                _workspace(dev).data_ptr(), st), "bn_bwd_reduce")
            dx = torch.empty_like(x, memory_format=torch.channels_last)
            _lib.check(lib.tfop_bn_bwd_apply(
                dz.data_ptr(), x.data_ptr(), 0, stats[0].data_ptr(), stats[1].data_ptr(),
                gamma.data_ptr(), pd, coef.data_ptr(), dx.data_ptr(), 0, m, c, st), "bn_bwd_apply")
            return dx, dgamma, dbeta, None, None, None, None
        _lib.check(lib.tfop_stem_bn_pool_bwd_reduce(
            dy.data_ptr(), idx.data_ptr(), x.data_ptr(), stats[0].data_ptr(), stats[1].data_ptr(),
            n, h, w, c, dgamma.data_ptr(), dbeta.data_ptr(), pd, coef.data_ptr(),
            _workspace(dev).data_ptr(), st), "stem_bn_pool_bwd_reduce")
        dx = torch.empty_like(x, memory_format=torch.channels_last)
        _lib.check(lib.tfop_stem_bn_pool_bwd_apply(
            dy.data_ptr(), idx.data_ptr(), x.data_ptr(), stats[0].data_ptr(), stats[1].data_ptr(),
            gamma.data_ptr(), pd, coef.data_ptr(), dx.data_ptr(), n, h, w, c, st),
            "stem_bn_pool_bwd_apply")
        return dx, dgamma, dbeta, None, None, None, None


def bn_relu_maxpool(bn: "FusedBatchNorm2d", x: torch.Tensor) -> torch.Tensor:
    """maxpool3x3/2/pad1(relu(bn(x))): one fused pass in training mode on CUDA bf16 NHWC,
    the two separate ops otherwise."""
    if bn.training and bn.relu and supported(x) and x.numel() // x.shape[1] < 2 ** 31:
        if bn.running_mean.dtype != torch.float32:
            bn.running_mean = bn.running_mean.float()
            bn.running_var = bn.running_var.float()
        return _BnReluMaxPool.apply(x, bn.weight, bn.bias, bn.running_mean, bn.running_var,
                                    bn.momentum, bn.eps)
    from .pool import max_pool_3x3_s2
    return max_pool_3x3_s2(bn(x))


class FusedBatchNorm2d(nn.Module):
    """BatchNorm2d with optional fused residual-add and ReLU (drop-in for BN+ReLU)."""

    def __init__(self, num_features: int, relu: bool = True, eps: float = 1e-5,
                 momentum: float = 0.1, zero_init: bool = False):
        super().__init__()
        self.num_features = num_features
        self.relu = relu
        self.eps = eps
        self.momentum = momentum
        self.weight = nn.Parameter(torch.zeros(num_features) if zero_init
                                   else torch.ones(num_features))
        self.bias = nn.Parameter(torch.zeros(num_features))
        self.register_buffer("running_mean", torch.zeros(num_features, dtype=torch.float32))
        self.register_buffer("running_var", torch.ones(num_features, dtype=torch.float32))

    def forward(self, x, residual=None, sums=None):
        if self.running_mean.dtype != torch.float32:
            # a blanket .to(bfloat16) also casts buffers; statistics stay fp32
            self.running_mean = self.running_mean.float()
            self.running_var = self.running_var.float()
        return fused_bn_act(x, self.weight, self.bias, self.running_mean, self.running_var,
                            self.training, self.momentum, self.eps, self.relu, residual, sums)
