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


def _operand(t: torch.Tensor) -> torch.Tensor:
    if t.stride(1) != 1 or t.stride(0) % 8 or t.data_ptr() % 16:
        t = t.contiguous()
        if t.shape[1] % 8:
            t = F.pad(t, (0, 8 - t.shape[1] % 8))
    return t


def gemm(a: torch.Tensor, b: torch.Tensor, bias: torch.Tensor | None = None, epilogue: str = "none",
         out_dtype=torch.bfloat16, tile_n: int = 0, a_t: bool = False, b_t: bool = False,
         out: torch.Tensor | None = None, stream=None) -> torch.Tensor:
    """General tcgen05 GEMM, C[M,N] = epi(op(a) @ op(b)^T + bias) with no transposed copies:

      a_t=False: a is [M, K];  a_t=True: a is stored [K, M] (the tensor core reads it MN-major)
      b_t=False: b is [N, K];  b_t=True: b is stored [K, N]

    ``gemm(x, W)`` = F.linear; ``gemm(dy, W, b_t=True)`` = dy @ W (dgrad);
    ``gemm(dy, x, a_t=True, b_t=True)`` = dy^T @ x (wgrad).
    """
    assert a.is_cuda and a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16
    assert a.dim() == 2 and b.dim() == 2
    m, k = (a.shape[1], a.shape[0]) if a_t else a.shape
    n, kb = (b.shape[1], b.shape[0]) if b_t else b.shape
    assert k == kb, f"contraction mismatch {k} vs {kb}"
    # row pitch must be a 16-byte multiple; padded columns lie outside the TMA tensor extents
    # (true M/N/K are passed), so they are never read
    a, b = _operand(a), _operand(b)
    if out is None:
        out = torch.empty((m, n), dtype=out_dtype, device=a.device)
    epi = EPI[epilogue]
    if bias is not None and epi == 0:
        epi = 1
    _lib.check(_lib.load().tfop_gemm_bf16(
        a.data_ptr(), b.data_ptr(), out.data_ptr(), _lib.ptr(bias),
        _lib.dtype_tag(bias.dtype) if bias is not None else 0, m, n, k, a.stride(0), b.stride(0),
        out.stride(0), _lib.dtype_tag(out.dtype), epi, tile_n, int(a_t), int(b_t),
        _lib.stream_handle(stream)), "gemm_bf16")
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
            # activation backward straight from the saved pre-activation (no forward re-run)
            if ctx.act == "relu":
                dy2 = torch.ops.aten.threshold_backward(dy2, pre, 0)
            else:
                dy2 = torch.ops.aten.gelu_backward(
                    dy2, pre, approximate="tanh" if ctx.act == "gelu_tanh" else "none")
        dy2 = dy2.contiguous()
        dx = dw = db = None
        mn_ok = dy2.shape[1] % 8 == 0 and w.shape[1] % 8 == 0 and x2.shape[1] % 8 == 0
        if ctx.needs_input_grad[0]:
            # dX[M,K] = dY[M,N] @ W[N,K]: W is the [contraction, K] matrix as stored -> MN-major B
            dx = (gemm(dy2, w, b_t=True) if mn_ok else gemm_tn(dy2, transpose(w))).view(ctx.xshape)
        if ctx.needs_input_grad[1]:
            # dW[N,K] = dY^T @ X: both stored [tokens, *] -> both MN-major, contraction over rows
            # a flat-buffer engine lets the GEMM store dW straight into the gradient bucket
            target = getattr(w, "_tfop_grad_target", None)
            out = target.take() if (target is not None and mn_ok and w.dtype == torch.bfloat16) else None
            if out is not None and target.fused is not None:
                # data-parallel job: dW is produced and summed over the ranks by one kernel,
                # in place in the multicast-bound gradient buffer (the engine's fused_wgrad)
                comm, ws_off, scale, base = target.fused
                comm.gemm_all_reduce_nvls(dy2, x2, a_mn=True, b_mn=True, ws_off=ws_off, scale=scale,
                                          out_off=out.data_ptr() - base)
                dw = out
            else:
                dw = (gemm(dy2, x2, a_t=True, b_t=True, out=out) if mn_ok
                      else gemm_tn(transpose(dy2), transpose(x2)))
        if ctx.has_bias and ctx.needs_input_grad[2]:
            from .layernorm import bias_grad
            db = bias_grad(dy2)
        return dx, dw, db, None


def linear(x, w, b=None, act: str = "none"):
    """act(x @ w.T + b): the tcgen05 kernel for CUDA bf16 tensors, plain torch on the CPU. CUDA
    tensors of another dtype raise - a layer built from this module never drops to a library
    GEMM behind the caller's back (cast the model to bf16, or build it from nn.Linear)."""
    if x.is_cuda:
        if x.dtype != torch.bfloat16 or w.dtype != torch.bfloat16:
            raise TypeError(f"tcgen05 linear needs bf16 CUDA tensors, got {x.dtype} / {w.dtype}")
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
