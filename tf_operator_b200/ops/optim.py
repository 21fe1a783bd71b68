"""Flat-buffer fused optimizers (csrc/optim.cu) and the multi-tensor gradient pack.

All parameters of a model live in four flat buffers (bf16 weights the kernels
read, fp32 master weights, fp32 state, bf16 gradient bucket). One optimizer
step is ONE kernel launch; hyper-parameters sit in device memory so a captured
CUDA graph follows the LR schedule. Optimizer step of the reference workloads:
examples/v1/dist-mnist/dist_mnist.py:194-208.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib

HP_LR, HP_MOM, HP_WD, HP_GSCALE, HP_B1, HP_B2, HP_EPS, HP_STEP = range(8)


def sgd_momentum_(master, mom, grad, weights_bf16, n_decay, hp, nesterov=False, stream=None):
    lib = _lib.load()
    _lib.check(lib.tfop_sgd_momentum(
        master.data_ptr(), mom.data_ptr(), grad.data_ptr(), _lib.dtype_tag(grad.dtype),
        _lib.ptr(weights_bf16), master.numel(), int(n_decay), hp.data_ptr(), int(nesterov),
        _lib.stream_handle(stream)), "sgd_momentum")


def adamw_(master, m1, m2, grad, weights_bf16, n_decay, hp, stream=None):
    lib = _lib.load()
    _lib.check(lib.tfop_adamw(
        master.data_ptr(), m1.data_ptr(), m2.data_ptr(), grad.data_ptr(),
        _lib.dtype_tag(grad.dtype), _lib.ptr(weights_bf16), master.numel(), int(n_decay),
        hp.data_ptr(), _lib.stream_handle(stream)), "adamw")


def zero_(t: torch.Tensor, stream=None):
    lib = _lib.load()
    _lib.check(lib.tfop_zero(t.data_ptr(), t.numel() * t.element_size(),
                             _lib.stream_handle(stream)), "zero")


def multi_pack_bf16(srcs, dst_offsets, dst: torch.Tensor, scale: float = 1.0, stream=None):
    """Copy+cast tensors ``srcs`` (all fp32 or all bf16, dense) into ``dst`` (flat bf16) at
    element offsets ``dst_offsets`` (multiples of 8). One launch per 128 tensors."""
    lib = _lib.load()
    if not srcs:
        return 0
    tag = _lib.dtype_tag(srcs[0].dtype)
    st = _lib.stream_handle(stream)
    launches = 0
    for i in range(0, len(srcs), 128):
        chunk = srcs[i:i + 128]
        segs = (_lib.PackSeg * len(chunk))()
        for j, (t, off) in enumerate(zip(chunk, dst_offsets[i:i + 128])):
            assert _lib.dtype_tag(t.dtype) == tag, "mixed gradient dtypes in one pack call"
            segs[j].src = t.data_ptr()
            segs[j].dst_off = off
            segs[j].n = t.numel()
        _lib.check(lib.tfop_multi_pack_bf16(C.cast(segs, C.c_void_p), len(chunk), tag,
                                            dst.data_ptr(), float(scale), st), "multi_pack")
        launches += 1
    return launches
