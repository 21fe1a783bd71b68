"""Multi-head self-attention on the tcgen05 kernels of csrc/attention.cu (head dimension 64,
no mask): ``attention(qkv, heads)`` takes the PACKED projection output ``[B, S, 3H]`` and returns
the context ``[B, S, H]``; the backward returns the gradient in the same packed layout. No
``unbind`` / ``transpose`` / ``cat`` copies on either side.

CPU tensors (and the library-twin model) use ``F.scaled_dot_product_attention``. CUDA tensors the
kernels cannot take raise instead of falling back.

Reference parity: the worker-side math of the BERT-large TFJob of BASELINE.json config 3
(examples/v1/dist-mnist/dist_mnist.py:188-192 is where the reference's example trainers compute).
"""
from __future__ import annotations

import torch
import torch.nn.functional as F

from . import _lib


def _check(qkv: torch.Tensor, heads: int):
    if qkv.dtype != torch.bfloat16:
        raise TypeError(f"tcgen05 attention needs bf16, got {qkv.dtype}")
    b, s, h3 = qkv.shape
    if h3 % (3 * heads) or h3 // (3 * heads) != 64:
        raise ValueError(f"tcgen05 attention covers head dimension 64, got {h3 // (3 * heads)}")
    if s % 128:
        raise ValueError(f"sequence length must be a multiple of 128, got {s}")
    return b, s, h3 // 3


class _AttnFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, qkv, heads):
        b, s, h = _check(qkv, heads)
        qkv = qkv.contiguous()
        out = torch.empty((b, s, h), dtype=qkv.dtype, device=qkv.device)
        lse = torch.empty((b, heads, s), dtype=torch.float32, device=qkv.device)
        _lib.check(_lib.load().tfop_attn_fwd(qkv.data_ptr(), out.data_ptr(), lse.data_ptr(), b, s,
                                             heads, 64, _lib.stream_handle()), "attn_fwd")
        ctx.save_for_backward(qkv, out, lse)
        ctx.heads = heads
        return out

    @staticmethod
    def backward(ctx, dout):
        qkv, out, lse = ctx.saved_tensors
        b, s, h3 = qkv.shape
        h, heads = h3 // 3, ctx.heads
        dout = dout.contiguous()
        dqkv = torch.empty_like(qkv)
        delta = torch.empty((b, heads, s), dtype=torch.float32, device=qkv.device)
        dq32 = torch.empty((b * s, h), dtype=torch.float32, device=qkv.device)
        _lib.check(_lib.load().tfop_attn_bwd(qkv.data_ptr(), out.data_ptr(), dout.data_ptr(),
                                             lse.data_ptr(), dqkv.data_ptr(), delta.data_ptr(),
                                             dq32.data_ptr(), b, s, heads, 64, _lib.stream_handle()),
                   "attn_bwd")
        return dqkv, None


def attention_reference(qkv: torch.Tensor, heads: int) -> torch.Tensor:
    """The same function from library ops (fp32-capable oracle; CPU path)."""
    b, s, h3 = qkv.shape
    q, k, v = qkv.view(b, s, 3, heads, h3 // (3 * heads)).unbind(2)
    a = F.scaled_dot_product_attention(q.transpose(1, 2), k.transpose(1, 2), v.transpose(1, 2))
    return a.transpose(1, 2).reshape(b, s, h3 // 3)


def attention(qkv: torch.Tensor, heads: int) -> torch.Tensor:
    """softmax(Q K^T / sqrt(d)) V for qkv = [Q | K | V] packed along the last dimension."""
    if qkv.is_cuda:
        return _AttnFn.apply(qkv, heads)
    return attention_reference(qkv, heads)
