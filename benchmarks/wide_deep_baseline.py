#!/usr/bin/env python
"""Comparator for BASELINE.json config 4 (wide-and-deep, 10M-row embedding sharded over PS
replicas): the same model, shapes and id distribution as examples/wide_deep_ps/train.py, written
the way a stock PyTorch + NCCL user would - no kernel or engine of this repository on the path.

  ranks [0, P)      parameter servers: own rows r with r % P == rank (deep table dim 64 and wide
                    table dim 4 side by side), answer lookups and apply Adagrad;
  ranks [P, P + W)  workers: route the batch's ids to the owning shards with
                    ``all_to_all_single`` (counts, then ids), receive the rows the same way,
                    run the dense tower (torch.nn.Linear in bf16, DistributedDataParallel over
                    the worker group, torch AdamW), send the row gradients back; the PS merges
                    duplicates (``torch.unique`` + ``index_add_``) and updates.

This is the pull / push of the reference's PS example
(examples/v1/dist-mnist/dist_mnist.py:161-208: variables placed on ps tasks by
replica_device_setter, RecvTensor pull, gradient push, optimizer on the PS) carried by the
library collective a torch user has, instead of TF's gRPC runtime (not installable here).

    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 \
        benchmarks/wide_deep_baseline.py --ps 2

prints one JSON line from rank P (first worker): samples/s summed over the workers.
"""
import argparse
import json
import os

import torch
import torch.distributed as dist
import torch.nn as nn
import torch.nn.functional as F


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ps", type=int, default=2)
    ap.add_argument("--rows", type=int, default=10_000_000)
    ap.add_argument("--dim", type=int, default=64)
    ap.add_argument("--fields", type=int, default=26)
    ap.add_argument("--batch-size", type=int, default=8192)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--out", default="")
    ap.add_argument("--cpu", action="store_true", help="gloo on CPU: a smoke test of the routing")
    args = ap.parse_args()

    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    if args.cpu:
        dev = torch.device("cpu")
        dist.init_process_group("gloo")
    else:
        torch.cuda.set_device(local)
        dev = torch.device("cuda", local)
        dist.init_process_group("nccl", device_id=dev)
    P, W = args.ps, world - args.ps
    assert P >= 1 and W >= 1, "need at least one PS and one worker"
    is_ps = rank < P
    workers = dist.new_group(list(range(P, world)))
    D = args.dim + 4                                  # deep row | wide row
    rps = (args.rows + P - 1) // P
    lr, eps = 0.05, 1e-10

    if is_ps:
        g = torch.Generator(device=dev).manual_seed(rank)
        weight = torch.empty(rps, D, device=dev).normal_(0.0, 0.01, generator=g)
        accum = torch.full((rps, D), 0.1, device=dev)
    else:
        torch.manual_seed(1234)
        in_f = (args.fields * args.dim + 16 + 7) // 8 * 8
        mod = Dense(in_f).to(dev)
        mod.tower.to(torch.bfloat16)
        ddp = nn.parallel.DistributedDataParallel(mod, device_ids=None if args.cpu else [local],
                                                  process_group=workers) \
            if W > 1 else mod
        opt = torch.optim.AdamW(mod.parameters(), lr=1e-3, weight_decay=0.0, fused=not args.cpu)
        gen = torch.Generator(device=dev).manual_seed(7 + rank - P)
        pad = in_f - (args.fields * args.dim + 16)
    B = args.batch_size

    def exchange_counts(send_counts):
        recv = torch.empty_like(send_counts)
        dist.all_to_all_single(recv, send_counts)
        return recv

    def step():
        # ---- pull --------------------------------------------------------------------------
        if is_ps:
            send_counts = torch.zeros(world, dtype=torch.int64, device=dev)
            ids_sorted = torch.empty(0, dtype=torch.int64, device=dev)
        else:
            u = torch.rand(B, args.fields, device=dev, generator=gen)
            ids = (u.pow(3.0) * (args.rows - 1)).long().view(-1)
            dense = torch.randn(B, 16, device=dev, generator=gen)
            y = (torch.rand(B, device=dev, generator=gen) < 0.25).float()
            shard = ids % P
            order = torch.argsort(shard, stable=True)
            ids_sorted = ids[order] // P
            send_counts = torch.zeros(world, dtype=torch.int64, device=dev)
            send_counts[:P] = torch.bincount(shard, minlength=P)
        recv_counts = exchange_counts(send_counts)
        sc, rc = send_counts.tolist(), recv_counts.tolist()          # host sync, as torch needs
        got_ids = torch.empty(sum(rc), dtype=torch.int64, device=dev)
        dist.all_to_all_single(got_ids, ids_sorted, rc, sc)
        rows_out = weight[got_ids] if is_ps else torch.empty(0, D, device=dev)
        rows_in = torch.empty(sum(sc), D, device=dev)
        dist.all_to_all_single(rows_in, rows_out, sc, rc)
        # ---- dense tower on the workers ----------------------------------------------------
        if is_ps:
            grad_rows = torch.empty(0, D, device=dev)
            loss = None
        else:
            rows = torch.empty_like(rows_in)
            rows[order] = rows_in
            rows.requires_grad_(True)
            r = rows.view(B, args.fields, D)
            x = torch.cat([r[..., : args.dim].reshape(B, -1).to(torch.bfloat16),
                           dense.to(torch.bfloat16)], dim=1)
            if pad:
                x = F.pad(x, (0, pad))
            logits = ddp(x, r[..., args.dim].sum(dim=1))
            loss = F.binary_cross_entropy_with_logits(logits, y)
            opt.zero_grad(set_to_none=True)
            loss.backward()
            opt.step()
            grad_rows = rows.grad[order].contiguous()
        # ---- push --------------------------------------------------------------------------
        got_grads = torch.empty(sum(rc), D, device=dev)
        dist.all_to_all_single(got_grads, grad_rows, rc, sc)
        if is_ps and got_ids.numel():
            uniq, inv = torch.unique(got_ids, return_inverse=True)
            gsum = torch.zeros(uniq.numel(), D, device=dev).index_add_(0, inv, got_grads)
            acc = accum[uniq] + gsum * gsum
            accum[uniq] = acc
            weight[uniq] -= lr * gsum / (acc.sqrt() + eps)
        return loss

    for _ in range(args.warmup):
        step()
    if args.cpu:
        import time
        dist.barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            loss = step()
        ms = torch.tensor([(time.perf_counter() - t0) * 1e3 / args.steps], dtype=torch.float64)
    else:
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.steps):
            loss = step()
        e1.record()
        torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1) / args.steps], device=dev, dtype=torch.float64)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    if rank == P:
        out = {"impl": "stock torch: NCCL all_to_all pull/push, torch.unique + index_add_ Adagrad "
                       "on the PS ranks, DDP + fused AdamW dense tower",
               "ps": P, "workers": W, "rows": args.rows, "dim": args.dim, "fields": args.fields,
               "batch_per_worker": B, "ms_per_step": round(float(ms.item()), 3),
               "samples_per_s": round(W * B / float(ms.item()) * 1e3, 1),
               "loss": round(float(loss.detach()), 4), "timing": "CUDA events, max over ranks"}
        line = json.dumps(out)
        print(line, flush=True)
        if args.out:
            with open(args.out, "w") as f:
                f.write(line + "\n")
    dist.barrier()
    dist.destroy_process_group()


class Dense(nn.Module):
    def __init__(self, in_f):
        super().__init__()
        self.tower = nn.Sequential(nn.Linear(in_f, 1024), nn.ReLU(), nn.Linear(1024, 512), nn.ReLU(),
                                   nn.Linear(512, 256), nn.ReLU(), nn.Linear(256, 1))
        self.wide_bias = nn.Parameter(torch.zeros(1))

    def forward(self, x, wide):
        return self.tower(x)[:, 0].float() + wide + self.wide_bias


if __name__ == "__main__":
    main()
