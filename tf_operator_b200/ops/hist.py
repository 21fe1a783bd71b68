"""Gradient histograms for histogram-based boosting (csrc/hist.cu) + their all-reduce (K6).

The XGBoostJob path of the reference (examples/xgboost/xgboostjob.yaml:19-25) builds
per-worker histograms and all-reduces them with rabit over TCP; here the build is a CUDA
kernel and the all-reduce is the fp32 P2P / NVLS kernel of csrc/allreduce.cu / nvls.cu.
"""
from __future__ import annotations

from typing import Optional

import torch

from . import _lib

NUM_BINS = 256


def build_histogram(bins_u8: torch.Tensor, grad: torch.Tensor, hess: torch.Tensor,
                    row_idx: Optional[torch.Tensor] = None, out: Optional[torch.Tensor] = None,
                    stream=None) -> torch.Tensor:
    """bins_u8: [rows, F] uint8; grad/hess: [rows] fp32; row_idx: int32 subset or None.
    Returns fp32 [F, 256, 2] with (sum grad, sum hess) per feature bin."""
    rows, f = bins_u8.shape
    if out is None:
        out = torch.zeros((f, NUM_BINS, 2), dtype=torch.float32, device=bins_u8.device)
    else:
        out.zero_()
    if not bins_u8.is_cuda:
        idx = row_idx.long() if row_idx is not None else torch.arange(rows)
        b = bins_u8[idx].long()
        for j in range(f):
            out[j, :, 0].index_add_(0, b[:, j], grad[idx])
            out[j, :, 1].index_add_(0, b[:, j], hess[idx])
        return out
    n = row_idx.numel() if row_idx is not None else rows
    _lib.check(_lib.load().tfop_hist_build(
        bins_u8.data_ptr(), grad.data_ptr(), hess.data_ptr(), _lib.ptr(row_idx), n, f,
        out.data_ptr(), _lib.stream_handle(stream)), "hist_build")
    return out
