"""tcgen05 / TMEM / TMA bf16 GEMM (csrc/gemm_tcgen05.cu) and a Linear layer on top of it.

``linear(x, w, b, act)`` == act(x @ w.T + b). Forward always runs the hand-written
tcgen05 kernel on CUDA; the backward GEMMs (dX = dY @ W, dW = dY^T @ X) are
expressed through the same TN kernel on transposed copies.
Dense layers of the reference trainers: examples/v1/dist-mnist/dist_mnist.py:188-192.
"""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import _lib

EPI = {"none": 0, "bias": 1, "gelu": 2, "relu": 3, "gelu_tanh": 4}


def gemm_tn(a: torch.Tensor, b: torch.Tensor, bias: torch.Tensor | None = None,
            epilogue: str = "none", out_dtype=torch.bfloat16, tile_n: int = 0,
            out: torch.Tensor | None = None, stream=None) -> torch.Tensor:
    """C[M,N] = epi(A[M,K] @ B[N,K]^T + bias). A, B: bf16, last dim contiguous."""
    assert a.is_cuda and a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16
    assert a.dim() == 2 and b.dim() == 2 and a.shape[1] == b.shape[1]
    if a.stride(1) != 1 or a.stride(0) % 8 or a.data_ptr() % 16:
        a = a.contiguous()
    if b.stride(1) != 1 or b.stride(0) % 8 or b.data_ptr() % 16:
        b = b.contiguous()
    m, k = a.shape
    n = b.shape[0]
    if k % 8:
        # TMA needs 16-byte row pitch: pad K (zeros do not change the product)
        a = F.pad(a, (0, 8 - k % 8))
        b = F.pad(b, (0, 8 - k % 8))
        k = a.shape[1]
    if out is None:
        out = torch.empty((m, n), dtype=out_dtype, device=a.device)
    epi = EPI[epilogue]
    if bias is not None and epi == 0:
        epi = 1
    lib = _lib.load()
    _lib.check(lib.tfop_gemm_bf16_tn(
        a.data_ptr(), b.data_ptr(), out.data_ptr(), _lib.ptr(bias),
        _lib.dtype_tag(bias.dtype) if bias is not None else 0, m, n, k, a.stride(0), b.stride(0),
        out.stride(0), _lib.dtype_tag(out.dtype), epi, tile_n, _lib.stream_handle(stream)),
        "gemm_bf16_tn")
    return out


def transpose(x: torch.Tensor) -> torch.Tensor:
    """x[R, C] -> contiguous [C, R]; CUDA bf16 uses csrc/transpose.cu."""
    if not (x.is_cuda and x.dtype == torch.bfloat16 and x.dim() == 2 and x.stride(1) == 1):
        return x.t().contiguous()
    r, c = x.shape
    out = torch.empty((c, r), dtype=x.dtype, device=x.device)
    _lib.check(_lib.load().tfop_transpose_bf16(x.data_ptr(), out.data_ptr(), r, c, x.stride(0), r,
                                               _lib.stream_handle()), "transpose_bf16")
    return out


class _LinearFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, b, act):
        x2 = x.reshape(-1, x.shape[-1])
        if act in ("gelu", "relu", "gelu_tanh") and any(ctx.needs_input_grad[:3]):
            pre = gemm_tn(x2, w, b, "bias" if b is not None else "none")
            y = {"gelu": lambda t: F.gelu(t), "relu": F.relu,
                 "gelu_tanh": lambda t: F.gelu(t, approximate="tanh")}[act](pre)
            ctx.save_for_backward(x2, w, pre)
        else:
            y = gemm_tn(x2, w, b, act if b is not None else "none")
            ctx.save_for_backward(x2, w, None)
        ctx.act = act
        ctx.has_bias = b is not None
        ctx.xshape = x.shape
        return y.view(*x.shape[:-1], w.shape[0])

    @staticmethod
    def backward(ctx, dy):
        x2, w, pre = ctx.saved_tensors
        dy2 = dy.reshape(-1, dy.shape[-1])
        if pre is not None:
            with torch.enable_grad():
                p = pre.detach().requires_grad_(True)
                a = {"gelu": lambda t: F.gelu(t), "relu": F.relu,
                     "gelu_tanh": lambda t: F.gelu(t, approximate="tanh")}[ctx.act](p)
            dy2, = torch.autograd.grad(a, p, dy2)
        dy2 = dy2.contiguous()
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            # dX[M,K] = dY[M,N] @ W[N,K]  ==  TN(dY, W^T[K,N])
            dx = gemm_tn(dy2, transpose(w)).view(ctx.xshape)
        if ctx.needs_input_grad[1]:
            # dW[N,K] = dY^T[N,M] @ X[M,K]  ==  TN(dY^T[N,M], X^T[K,M])
            dw = gemm_tn(transpose(dy2), transpose(x2))
        if ctx.has_bias and ctx.needs_input_grad[2]:
            db = dy2.sum(0)
        return dx, dw, db, None


def linear(x, w, b=None, act: str = "none"):
    """act(x @ w.T + b) — tcgen05 kernel for CUDA bf16, torch for everything else."""
    if x.is_cuda and x.dtype == torch.bfloat16 and w.dtype == torch.bfloat16:
        return _LinearFn.apply(x, w, b, act)
    y = F.linear(x, w, b)
    if act == "gelu":
        y = F.gelu(y)
    elif act == "gelu_tanh":
        y = F.gelu(y, approximate="tanh")
    elif act == "relu":
        y = F.relu(y)
    return y


class TcLinear(nn.Linear):
    """nn.Linear whose CUDA bf16 path is the hand-written tcgen05 GEMM."""

    def __init__(self, in_features, out_features, bias=True, act: str = "none"):
        super().__init__(in_features, out_features, bias=bias)
        self.act = act

    def forward(self, x):
        return linear(x, self.weight, self.bias, self.act)
