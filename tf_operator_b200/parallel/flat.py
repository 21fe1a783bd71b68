"""Flat parameter / gradient / optimizer-state buffers.

Every parameter of the model becomes a strided view into one flat bf16 weight
buffer; gradients are packed (cast fused) into one flat bf16 bucket buffer —
placed in symmetric peer memory when a communicator is given, so the
all-reduce runs on it in place with zero copies — and the optimizer is a single
fused kernel over the flat fp32 master / state buffers.

Layout: [ weight-decay params (reverse registration order) | no-decay params ],
each slot padded to 8 elements (16 B in bf16). Reverse order means the slots
fill front-to-back as backward proceeds, so buckets are contiguous ranges.

Counterpart in the reference: none (the gradient exchange is a library call in
examples/v1/distribution_strategy/keras-API/multi_worker_strategy-with-keras.py:76-77).
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import List, Optional, Sequence

import torch
import torch.nn as nn


def _pad8(n: int) -> int:
    return (n + 7) // 8 * 8


@dataclass
class Slot:
    name: str
    param: nn.Parameter
    offset: int          # element offset into the flat buffers
    numel: int
    decay: bool


@dataclass
class Bucket:
    lo: int              # element range [lo, hi) in the flat buffers
    hi: int
    slots: List[Slot] = field(default_factory=list)
    pending: int = 0


class GradTarget:
    """Lets a producer write a parameter's gradient straight into its slot of the flat gradient
    bucket (ops/gemm._LinearFn does, for weight gradients): ``take()`` hands out a fresh view of
    the slot once per step - a second use of the same parameter in one step gets None and
    produces an ordinary tensor that autograd adds in place. Autograd adopts the returned view
    as ``param.grad`` without copying, and the engine's pack step skips gradients that already
    live in the bucket."""

    def __init__(self, flat: "FlatParams", slot: "Slot"):
        self.flat, self.slot, self.used_step = flat, slot, -1
        # (comm, ws_off, scale) when this parameter's weight gradient is produced AND summed over
        # the ranks by one kernel (SymmetricComm.gemm_all_reduce_nvls) straight into its slot
        self.fused = None

    def take(self) -> Optional[torch.Tensor]:
        if self.used_step == self.flat.step_id:
            return None
        self.used_step = self.flat.step_id
        p = self.slot.param
        return torch.as_strided(self.flat.grads, p.shape, p.stride(), self.slot.offset)


class FlatParams:
    def __init__(self, model: nn.Module, device: torch.device, *, grad_buffer: Optional[torch.Tensor] = None,
                 bucket_bytes: int = 16 << 20, weight_dtype: torch.dtype = torch.bfloat16,
                 fused: Optional[Sequence[nn.Parameter]] = None):
        """``fused``: parameters whose gradient arrives already summed over the ranks (see
        GradTarget.fused). They get the front of the flat buffers and belong to no bucket."""
        named = [(n, p) for n, p in model.named_parameters() if p.requires_grad]
        fused_ids = {id(p) for p in (fused or [])}
        decay = [(n, p) for n, p in named if p.ndim > 1 and id(p) not in fused_ids]
        fused_named = [(n, p) for n, p in named if id(p) in fused_ids]
        assert all(p.ndim > 1 for _, p in fused_named), "only weight matrices can be fused"
        nodecay = [(n, p) for n, p in named if p.ndim <= 1]
        self.slots: List[Slot] = []
        off = 0
        self.n_fused = 0
        for gi, (group, is_decay) in enumerate(((fused_named[::-1], True), (decay[::-1], True),
                                                (nodecay[::-1], False))):
            for n, p in group:
                self.slots.append(Slot(n, p, off, p.numel(), is_decay))
                off += _pad8(p.numel())
            if gi == 0:
                self.n_fused = off
            if is_decay:
                self.n_decay = off
        self.total = off
        self.device = device
        self.weight_dtype = weight_dtype
        self.master = torch.zeros(self.total, dtype=torch.float32, device=device)
        self.weights = torch.zeros(self.total, dtype=weight_dtype, device=device)
        if grad_buffer is not None:
            assert grad_buffer.numel() >= self.total and grad_buffer.dtype == torch.bfloat16
            self.grads = grad_buffer[: self.total]
        else:
            self.grads = torch.zeros(self.total, dtype=torch.bfloat16, device=device)
        # move parameter storage into the flat weight buffer, keeping each
        # parameter's shape AND strides (channels-last conv weights stay channels-last)
        with torch.no_grad():
            for s in self.slots:
                p = s.param
                src = p.detach().to(device)
                if not _is_dense(src):
                    src = src.contiguous()
                view = torch.as_strided(self.weights, src.shape, src.stride(), s.offset)
                mview = torch.as_strided(self.master, src.shape, src.stride(), s.offset)
                mview.copy_(src.float())
                view.copy_(src.to(weight_dtype))
                p.data = view
        # buckets: contiguous ranges, cut every bucket_bytes over the decay region. The no-decay
        # tail (norm scales / biases of EVERY layer, a few hundred KB) rides in the last decay
        # bucket: both become complete only when backward reaches the first layer, and one
        # latency-bound collective at the end of the step is cheaper than two.
        self.buckets: List[Bucket] = []
        cur = None
        limit = max(1, bucket_bytes // 2)
        for s in self.slots:
            if s.offset < self.n_fused:
                continue                      # reduced by the kernel that produces it
            if cur is None or (s.decay and cur.hi - cur.lo >= limit):
                cur = Bucket(s.offset, s.offset)
                self.buckets.append(cur)
            cur.slots.append(s)
            cur.hi = s.offset + _pad8(s.numel)
        # direct gradient targets (see GradTarget); step_id is bumped by the engine every step
        self.step_id = 0
        for s in self.slots:
            s.param._tfop_grad_target = GradTarget(self, s)
        self._bucket_of = {}
        for b in self.buckets:
            for s in b.slots:
                self._bucket_of[id(s.param)] = b

    def grad_in_place(self, s: Slot) -> bool:
        """True when the parameter's .grad already is its slot of the flat gradient buffer."""
        g = s.param.grad
        return g is not None and g.data_ptr() == self.grads.data_ptr() + s.offset * self.grads.element_size()

    def bucket_of(self, p: nn.Parameter) -> Optional[Bucket]:
        """None for a parameter of the fused region."""
        return self._bucket_of.get(id(p))

    def grad_view(self, s: Slot) -> torch.Tensor:
        p = s.param
        return torch.as_strided(self.grads, p.shape, p.stride(), s.offset)

    # ---- checkpoint support: flat buffers <-> per-parameter tensors keyed by name ----------
    def view_of(self, flat: torch.Tensor, s: Slot) -> torch.Tensor:
        """The slot's region of a flat buffer, shaped and strided like the parameter."""
        p = s.param
        return torch.as_strided(flat, p.shape, p.stride(), s.offset)

    def export_named(self, flat: torch.Tensor) -> dict:
        """{parameter name: contiguous CPU copy} of one flat buffer (master weights, momentum...).
        Layout-independent, so a checkpoint survives a change of bucket size or parameter order."""
        return {s.name: self.view_of(flat, s).detach().to("cpu", copy=True).contiguous()
                for s in self.slots}

    def import_named(self, flat: torch.Tensor, named: dict, strict: bool = True) -> List[str]:
        """Copy per-parameter tensors back into a flat buffer; returns the names not found."""
        missing = []
        with torch.no_grad():
            for s in self.slots:
                if s.name not in named:
                    missing.append(s.name)
                    continue
                src = named[s.name]
                if tuple(src.shape) != tuple(s.param.shape):
                    raise ValueError(f"checkpoint shape {tuple(src.shape)} != parameter shape "
                                     f"{tuple(s.param.shape)} for {s.name}")
                self.view_of(flat, s).copy_(src.to(flat.device, dtype=flat.dtype))
        if strict and missing:
            raise KeyError(f"checkpoint lacks parameters: {missing[:5]}{'...' if len(missing) > 5 else ''}")
        return missing

    def refresh_weights(self) -> None:
        """Model-dtype weights := round(master) after the master buffer was loaded."""
        with torch.no_grad():
            self.weights.copy_(self.master.to(self.weight_dtype))

    def check_grad_layouts(self) -> List[str]:
        """Names of parameters whose .grad strides differ from the parameter's."""
        def sig(t):
            return [(sz, st) for sz, st in zip(t.shape, t.stride()) if sz > 1]
        return [s.name for s in self.slots
                if s.param.grad is not None and sig(s.param.grad) != sig(s.param)]


def _is_dense(t: torch.Tensor) -> bool:
    """True when the tensor covers numel() distinct consecutive elements."""
    if t.numel() == 0:
        return True
    sizes_strides = sorted(((st, sz) for sz, st in zip(t.shape, t.stride()) if sz > 1))
    expect = 1
    for st, sz in sizes_strides:
        if st != expect:
            return False
        expect *= sz
    return True
