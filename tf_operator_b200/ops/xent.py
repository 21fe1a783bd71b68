"""Softmax cross-entropy with ignore_index over bf16 logits (csrc/xent.cu, K10).

``softmax_cross_entropy(logits, labels, ignore_index)`` == ``F.cross_entropy(logits.float(), labels,
ignore_index=...)`` (mean over the labelled rows) without the fp32 copy of the logits: forward
reads only labelled rows, backward writes bf16 dlogits directly. CUDA bf16 runs the kernels;
everything else runs the torch composition. No reference counterpart (BASELINE.json config 3).
"""
from __future__ import annotations

import torch
import torch.nn.functional as F

from . import _lib


class _SoftmaxXent(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, labels, ignore_index):
        rows, V = logits.shape
        labels = labels.contiguous()
        n_valid = (labels != ignore_index).sum()
        inv_n = (1.0 / n_valid.clamp(min=1).float()).reshape(1)
        stats = torch.empty((2, rows), dtype=torch.float32, device=logits.device)
        _lib.check(_lib.load().tfop_xent_fwd(
            logits.data_ptr(), logits.stride(0), labels.data_ptr(), rows, V, int(ignore_index),
            inv_n.data_ptr(), stats[0].data_ptr(), stats[1].data_ptr(), _lib.stream_handle()),
            "xent_fwd")
        ctx.save_for_backward(logits, labels, stats, inv_n)
        ctx.ignore_index = int(ignore_index)
        return stats[0].sum()

    @staticmethod
    def backward(ctx, grad_out):
        logits, labels, stats, inv_n = ctx.saved_tensors
        rows, V = logits.shape
        g = grad_out.detach().to(torch.float32).reshape(1).contiguous()
        dlogits = torch.empty((rows, V), dtype=logits.dtype, device=logits.device)
        _lib.check(_lib.load().tfop_xent_bwd(
            logits.data_ptr(), logits.stride(0), labels.data_ptr(), rows, V, ctx.ignore_index,
            inv_n.data_ptr(), stats[1].data_ptr(), g.data_ptr(), dlogits.data_ptr(),
            dlogits.stride(0), _lib.stream_handle()), "xent_bwd")
        return dlogits, None, None


def softmax_cross_entropy(logits: torch.Tensor, labels: torch.Tensor, ignore_index: int = -100):
    """Mean cross-entropy over rows whose label != ignore_index; logits [rows, V]."""
    if (logits.is_cuda and logits.dtype == torch.bfloat16 and logits.dim() == 2
            and logits.stride(1) == 1 and labels.dtype == torch.int64):
        return _SoftmaxXent.apply(logits, labels, ignore_index)
    return F.cross_entropy(logits.float(), labels, ignore_index=ignore_index)
