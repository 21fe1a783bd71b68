"""Sharded parameter-server embedding over NVLink peer memory (csrc/ps.cu, K4 + K5).

``ShardedEmbedding`` is the worker-side handle: ``lookup(ids)`` gathers rows straight out of
the owning PS shards' HBM; its backward segment-sums duplicate ids and applies Adagrad on the
owning shard with vector atomics over NVLink — the optimizer lives on the PS, workers never
hold embedding gradients or state (async PS training as in the reference's
examples/v1/dist-mnist/dist_mnist.py:161-208, on BASELINE.json config 4's shapes).
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional, Sequence

import torch

from . import _lib


class ShardedEmbedding:
    def __init__(self, weight_ptrs: Sequence[int], accum_ptrs: Sequence[int], rows_per_shard: int,
                 dim: int, lr: float = 0.05, eps: float = 1e-10, initial_accumulator: float = 0.1,
                 keepalive=None):
        assert 1 <= len(weight_ptrs) <= 8 and dim % 4 == 0
        self.table = _lib.PsTable()
        for i, (w, a) in enumerate(zip(weight_ptrs, accum_ptrs)):
            self.table.weight[i] = w
            self.table.accum[i] = a
        self.table.num_shards = len(weight_ptrs)
        self.table.dim = dim
        self.table.rows_per_shard = rows_per_shard
        self.num_shards, self.dim, self.rows_per_shard = len(weight_ptrs), dim, rows_per_shard
        self.lr, self.eps = lr, eps
        self.initial_accumulator = initial_accumulator
        self._keepalive = keepalive
        self.max_blocks = 0
        self._dedup_ws = {}          # (device index, positions) -> scratch of the device-side merge

    @property
    def num_rows(self) -> int:
        return self.rows_per_shard * self.num_shards

    # ---- construction helpers ---------------------------------------------------
    @classmethod
    def local(cls, num_rows: int, dim: int, num_shards: int = 1, device=None, init_std: float = 0.01,
              **kw) -> "ShardedEmbedding":
        """All shards in this process (tests, single-GPU runs)."""
        device = device or torch.device("cuda", torch.cuda.current_device())
        rps = (num_rows + num_shards - 1) // num_shards
        ws = [torch.randn(rps, dim, device=device) * init_std for _ in range(num_shards)]
        acc = [torch.full((rps, dim), kw.get("initial_accumulator", 0.1), device=device)
               for _ in range(num_shards)]
        emb = cls([w.data_ptr() for w in ws], [a.data_ptr() for a in acc], rps, dim,
                  keepalive=(ws, acc), **kw)
        emb.shards, emb.accums = ws, acc
        return emb

    @staticmethod
    def shard_bytes(rows_per_shard: int, dim: int) -> int:
        """Symmetric-buffer bytes a PS replica needs for its shard (weights + accumulators)."""
        return 2 * rows_per_shard * dim * 4

    @classmethod
    def from_comm(cls, comm, ps_ranks: Sequence[int], rows_per_shard: int, dim: int, base_off: int = 0,
                  **kw) -> "ShardedEmbedding":
        """Shards live at ``base_off`` in the symmetric buffers of ranks ``ps_ranks``."""
        n = rows_per_shard * dim * 4
        wp = [comm._lib.tfop_comm_buffer(comm._h, r) + base_off for r in ps_ranks]
        ap = [p + n for p in wp]
        return cls(wp, ap, rows_per_shard, dim, keepalive=comm, **kw)

    @staticmethod
    def init_local_shard(comm, rows_per_shard: int, dim: int, base_off: int = 0, init_std: float = 0.01,
                         initial_accumulator: float = 0.1, seed: int = 0) -> None:
        """Called by a PS replica: fill its own shard."""
        n = rows_per_shard * dim
        w = comm.view(base_off, n, torch.float32)
        a = comm.view(base_off + n * 4, n, torch.float32)
        g = torch.Generator(device=w.device).manual_seed(seed)
        w.normal_(0.0, init_std, generator=g)
        a.fill_(initial_accumulator)

    # ---- ops ------------------------------------------------------------------------
    def gather(self, ids: torch.Tensor, out_dtype=torch.float32, stream=None) -> torch.Tensor:
        ids = ids.contiguous().view(-1).to(torch.int64)
        out = torch.empty((ids.numel(), self.dim), dtype=out_dtype, device=ids.device)
        _lib.check(_lib.load().tfop_ps_gather(
            C.byref(self.table), ids.data_ptr(), ids.numel(), out.data_ptr(),
            _lib.dtype_tag(out_dtype), self.max_blocks, _lib.stream_handle(stream)), "ps_gather")
        return out

    def _workspace(self, device, n: int):
        key = (device.index, n)
        ws = self._dedup_ws.get(key)
        if ws is None:
            sizes = (C.c_size_t * 6)()
            _lib.check(_lib.load().tfop_ps_dedup_workspace(n, self.dim, sizes))
            ws = [torch.zeros(int(b), dtype=torch.uint8, device=device) for b in sizes]
            ws[0].fill_(0xFF)                   # empty key = all ones
            self._dedup_ws[key] = ws
        return ws

    def push(self, ids: torch.Tensor, grad: torch.Tensor, grad_scale: float = 1.0, stream=None) -> None:
        """Sum the gradients of duplicate ids and apply Adagrad on the owning shards: four kernels
        of csrc/ps.cu (K5b: hash insert, compact, sum, apply), no sort, no host round trip; the
        scratch is allocated once per batch shape."""
        ids = ids.contiguous().view(-1).to(torch.int64)
        grad = grad.contiguous().view(ids.numel(), self.dim)
        ws = self._workspace(ids.device, ids.numel())
        _lib.check(_lib.load().tfop_ps_push_adagrad_dedup(
            C.byref(self.table), ids.data_ptr(), ids.numel(), grad.data_ptr(),
            _lib.dtype_tag(grad.dtype), float(self.lr), float(self.eps), float(grad_scale),
            *[w.data_ptr() for w in ws], self.max_blocks, _lib.stream_handle(stream)),
            "ps_push_adagrad_dedup")
        _lib.LAUNCHES["ps_push_adagrad_dedup"] += 3       # four kernels behind the one entry point

    def push_sorted(self, ids: torch.Tensor, grad: torch.Tensor, grad_scale: float = 1.0, stream=None) -> None:
        """The same update with the duplicate merge done by a library sort (torch.sort +
        unique_consecutive + cumsum, then K5): the comparator of benchmarks/ps_bench.py."""
        ids = ids.contiguous().view(-1).to(torch.int64)
        grad = grad.contiguous().view(ids.numel(), self.dim)
        sorted_ids, perm = torch.sort(ids)
        uniq, counts = torch.unique_consecutive(sorted_ids, return_counts=True)
        seg = torch.zeros(uniq.numel() + 1, dtype=torch.int64, device=ids.device)
        torch.cumsum(counts, 0, out=seg[1:])
        _lib.check(_lib.load().tfop_ps_push_adagrad(
            C.byref(self.table), uniq.data_ptr(), seg.data_ptr(), perm.data_ptr(), uniq.numel(),
            grad.data_ptr(), _lib.dtype_tag(grad.dtype), float(self.lr), float(self.eps),
            float(grad_scale), self.max_blocks, _lib.stream_handle(stream)), "ps_push_adagrad")

    def lookup(self, ids: torch.Tensor, out_dtype=torch.float32) -> torch.Tensor:
        """Differentiable gather: backward pushes the gradient to the PS (returns [*ids.shape, dim])."""
        flat = _PsLookup.apply(self, ids, out_dtype, torch.empty(0, requires_grad=True, device=ids.device))
        return flat.view(*ids.shape, self.dim)


class _PsLookup(torch.autograd.Function):
    @staticmethod
    def forward(ctx, emb: ShardedEmbedding, ids, out_dtype, _hook):
        ctx.emb, ctx.ids = emb, ids
        return emb.gather(ids, out_dtype)

    @staticmethod
    def backward(ctx, dy):
        ctx.emb.push(ctx.ids, dy)
        return None, None, None, None
