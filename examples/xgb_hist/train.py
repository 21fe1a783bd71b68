#!/usr/bin/env python
"""Distributed histogram gradient boosting as an XGBoostJob replica (BASELINE.json config 5:
8 workers, 256 bins x 500 features, histogram all-reduce over NVLink).

Rendezvous follows the reference's env contract for XGBoost / LightGBM
(MASTER_ADDR / MASTER_PORT / WORLD_SIZE / RANK [+ WORKER_ADDRS], pkg/controller.v1/xgboost/
xgboost.go:59-94). Every replica holds a shard of the rows; per tree level it builds the
(grad, hess) histograms of the level's nodes locally (csrc/hist.cu), all-reduces them (fp32,
csrc/allreduce.cu one-shot / NVLS — the rabit all-reduce of the reference stack), and every
replica then picks the same best splits.
"""
import argparse
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))

from tf_operator_b200.ops.hist import NUM_BINS, build_histogram  # noqa: E402
from tf_operator_b200.runtime.env import ReplicaContext  # noqa: E402


def best_splits(hist, lam=1.0, min_hess=1e-3):
    """hist: [nodes, F, 256, 2] -> (gain, feature, bin) per node (exact greedy over bins)."""
    g, h = hist[..., 0], hist[..., 1]
    gl, hl = g.cumsum(-1), h.cumsum(-1)
    gt, ht = gl[..., -1:], hl[..., -1:]
    gr, hr = gt - gl, ht - hl
    gain = gl * gl / (hl + lam) + gr * gr / (hr + lam) - gt * gt / (ht + lam)
    gain = torch.where((hl > min_hess) & (hr > min_hess), gain, torch.full_like(gain, -1.0))
    flat = gain.flatten(1)
    best, arg = flat.max(dim=1)
    return best, arg // NUM_BINS, arg % NUM_BINS


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=1_000_000, help="rows per worker")
    ap.add_argument("--features", type=int, default=500)
    ap.add_argument("--trees", type=int, default=5)
    ap.add_argument("--depth", type=int, default=4)
    ap.add_argument("--lr", type=float, default=0.3)
    ap.add_argument("--impl", default="native", choices=["native", "nccl"])
    args = ap.parse_args()

    ctx = ReplicaContext.from_env()
    torch.cuda.set_device(0)
    dev = torch.device("cuda", 0)
    world, rank = ctx.world_size, ctx.rank
    hist_n = args.features * NUM_BINS * 2
    max_nodes = 2 ** (args.depth - 1)
    comm = None
    if world > 1:
        if args.impl == "nccl":
            ctx.init_process_group("nccl")
        else:
            comm = ctx.symmetric_comm(2 * max_nodes * hist_n * 4 + (2 << 20),
                                      nvls_bytes=max_nodes * hist_n * 4)

    g = torch.Generator(device=dev).manual_seed(100 + rank)
    bins = torch.randint(0, NUM_BINS, (args.rows, args.features), dtype=torch.uint8, device=dev, generator=g)
    w_true = torch.randn(args.features, device=dev, generator=torch.Generator(device=dev).manual_seed(1))
    y = ((bins[:, :32].float() - 127.5) @ w_true[:32] + 40 * torch.randn(args.rows, device=dev, generator=g) > 0).float()
    pred = torch.zeros(args.rows, device=dev)
    sync_buf = torch.zeros(8, device=dev)
    ar_us, ar_calls = 0.0, 0
    by_level = {}
    t0 = time.time()
    for tree in range(args.trees):
        p = torch.sigmoid(pred)
        grad, hess = p - y, (p * (1 - p)).clamp_min(1e-6)
        node = torch.zeros(args.rows, dtype=torch.int64, device=dev)
        leaf_value = torch.zeros(2 ** args.depth, device=dev)
        for level in range(args.depth):
            n_nodes = 2 ** level
            if comm is not None and comm.nvls_bytes:
                # build straight into the multicast-bound buffer: the all-reduce is then in place
                hist = comm.nvls_view(n_nodes * hist_n, torch.float32).view(
                    n_nodes, args.features, NUM_BINS, 2)
            else:
                hist = torch.zeros((n_nodes, args.features, NUM_BINS, 2), device=dev)
            for k in range(n_nodes):
                idx = (node == k).nonzero().flatten().int()
                build_histogram(bins, grad, hess, idx, out=hist[k])
            if world > 1:
                # line the ranks up first so the timing below is the all-reduce, not the skew
                # between the ranks' histogram builds
                if comm is not None:
                    comm.all_reduce(8, inp=sync_buf, out=sync_buf, wire_dtype=torch.float32)
                else:
                    torch.distributed.all_reduce(sync_buf)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                flat = hist.view(-1)
                if comm is not None and comm.nvls_bytes:
                    comm.all_reduce_nvls(flat.numel(), wire_dtype=torch.float32)
                elif comm is not None:
                    comm.all_reduce(flat.numel(), inp=flat, out=flat, wire_dtype=torch.float32,
                                    sym_off=1 << 20)
                else:
                    torch.distributed.all_reduce(flat)
                e1.record()
                torch.cuda.synchronize()
                dt_us = e0.elapsed_time(e1) * 1e3
                if tree > 0:                      # tree 0 carries one-time set-up costs
                    ar_us += dt_us
                    ar_calls += 1
                    by_level.setdefault(n_nodes, []).append(dt_us)
            if level == args.depth - 1:
                tot = hist[:, 0].sum(1)                       # [nodes, 2]
                leaf_value[:n_nodes] = -tot[:, 0] / (tot[:, 1] + 1.0)
                break
            _, feat, thr = best_splits(hist.clone() if (comm is not None and comm.nvls_bytes) else hist)
            go_right = bins[torch.arange(args.rows, device=dev), feat[node]].long() > thr[node]
            node = node * 2 + go_right.long()
        # leaves were indexed level by level: node ids at the last level
        pred += args.lr * leaf_value[node]
    torch.cuda.synchronize()
    dt = time.time() - t0
    p = torch.sigmoid(pred)
    loss = torch.nn.functional.binary_cross_entropy(p, y).item()
    acc = ((p > 0.5).float() == y).float().mean().item()
    out = {"rank": rank, "world": world, "impl": args.impl, "rows_per_worker": args.rows,
           "features": args.features, "trees": args.trees, "depth": args.depth,
           "logloss": round(loss, 4), "accuracy": round(acc, 4), "wall_s": round(dt, 3),
           "hist_allreduce_calls": ar_calls,
           "hist_allreduce_us_mean": round(ar_us / ar_calls, 1) if ar_calls else None,
           "hist_allreduce_us_by_level": {str(k): round(sum(v) / len(v), 1) for k, v in by_level.items()},
           "nvls": bool(comm is not None and comm.nvls_bytes)}
    print(json.dumps(out), flush=True)
    if comm is not None:
        time.sleep(0.5)
    return 0 if acc > 0.6 else 3


if __name__ == "__main__":
    sys.exit(main())
