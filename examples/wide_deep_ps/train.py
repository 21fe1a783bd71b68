#!/usr/bin/env python
"""Wide-and-deep with a parameter-server-sharded embedding (BASELINE.json config 4:
TFJob 2 PS + 6 Workers, 10M-row embedding, PS push/pull over NVLink P2P).

Roles come from TF_CONFIG exactly as in the reference's PS example
(examples/v1/dist-mnist/dist_mnist.py:102-146): a ``ps`` task owns one shard of the table in
its GPU's HBM and then just stays alive (``server.join()``); a ``worker`` task gathers the rows
it needs straight from the owning shards (K4) and, in backward, applies Adagrad on the owning
shard (K5). The dense tower (tcgen05 GEMMs) is replicated on the workers and all-reduced among
them with the fused P2P all-reduce.
"""
import argparse
import json
import os
import sys
import time

import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))

from tf_operator_b200.models.wide_deep import WideAndDeep  # noqa: E402
from tf_operator_b200.ops.ps_embedding import ShardedEmbedding  # noqa: E402
from tf_operator_b200.parallel.engine import FlatDataParallel  # noqa: E402
from tf_operator_b200.runtime.env import ReplicaContext, rendezvous_comm  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=10_000_000)
    ap.add_argument("--dim", type=int, default=64)
    ap.add_argument("--fields", type=int, default=26)
    ap.add_argument("--batch-size", type=int, default=8192)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--profile", action="store_true", help="print the step's kernels (torch profiler)")
    args = ap.parse_args()

    ctx = ReplicaContext.from_env()
    torch.cuda.set_device(0)
    dev = torch.device("cuda", 0)
    n_ps = len(ctx.cluster.get("ps", [])) or 1
    n_workers = len(ctx.cluster.get("worker", [])) or 1
    local = ctx.kind != "TFJob"
    rows_per_shard = (args.rows + n_ps - 1) // n_ps
    deep_bytes = ShardedEmbedding.shard_bytes(rows_per_shard, args.dim)
    wide_bytes = ShardedEmbedding.shard_bytes(rows_per_shard, 4)
    wide_off = (deep_bytes + 255) // 256 * 256

    if local:
        deep = ShardedEmbedding.local(args.rows, args.dim, 1, dev, lr=0.05)
        wide = ShardedEmbedding.local(args.rows, 4, 1, dev, lr=0.05)
        comm_w, rank_w = None, 0
    else:
        is_ps = ctx.task_type == "ps"
        rank = ctx.task_index if is_ps else n_ps + ctx.task_index
        world = n_ps + n_workers
        # peer handles go through the launcher's rendezvous store under <ns>/<job>/<job uid>/...
        # (a re-created job of the same name never sees its predecessor's handles) and in the
        # scope's current generation (a restarted replica re-forms the group)
        addr, uds = ctx.raw["TFOP_RENDEZVOUS"], ctx.raw.get("TFOP_RENDEZVOUS_UDS", "")
        nbytes = wide_off + wide_bytes if is_ps else (1 << 20)
        comm = rendezvous_comm(addr, uds, ctx.rendezvous_scope("ps"), rank, world, nbytes)
        rdv, prefix = comm._rdv
        if is_ps:
            ShardedEmbedding.init_local_shard(comm, rows_per_shard, args.dim, 0, seed=ctx.task_index)
            ShardedEmbedding.init_local_shard(comm, rows_per_shard, 4, wide_off, seed=100 + ctx.task_index)
            torch.cuda.synchronize()
            print(f"PS {ctx.task_index}: shard of {rows_per_shard} rows x ({args.dim}+4) ready "
                  f"({nbytes / 2**30:.2f} GiB)", flush=True)
        rdv.barrier(prefix + "/shards-ready", world)
        if is_ps:
            # server.join(): the shard lives in this process's HBM until the job is cleaned up
            # (cleanPodPolicy Running). Nothing runs here: the workers read and update the shard
            # over NVLink, and an idle owner serves them at full speed (benchmarks/ps_bench.py).
            while True:
                time.sleep(3600)
        ps_ranks = list(range(n_ps))
        deep = ShardedEmbedding.from_comm(comm, ps_ranks, rows_per_shard, args.dim, 0, lr=0.05)
        wide = ShardedEmbedding.from_comm(comm, ps_ranks, rows_per_shard, 4, wide_off, lr=0.05)
        rank_w = ctx.task_index
        comm_w = None
        if n_workers > 1:
            comm_w = rendezvous_comm(addr, uds, ctx.rendezvous_scope("workers"), rank_w, n_workers,
                                     64 << 20)

    torch.manual_seed(1234)
    model = WideAndDeep(deep, wide, num_fields=args.fields, embed_dim=args.dim).to(dev).to(torch.bfloat16)
    dp = FlatDataParallel(model, comm=comm_w, optimizer="adamw", lr=1e-3, weight_decay=0.0)
    g = torch.Generator(device=dev).manual_seed(7 + rank_w)
    B = args.batch_size

    def batch():
        # zipf-like id distribution over the 10M rows, dense features, clicks
        u = torch.rand(B, args.fields, device=dev, generator=g)
        ids = (u.pow(3.0) * (args.rows - 1)).long()
        dense = torch.randn(B, 16, device=dev, generator=g)
        y = (torch.rand(B, device=dev, generator=g) < 0.25).float()
        return ids, dense, y

    def step():
        ids, dense, y = batch()
        return dp.train_step(lambda: F.binary_cross_entropy_with_logits(model(ids, dense), y))

    for _ in range(args.warmup):
        step()
    torch.cuda.synchronize()
    if args.profile:
        import subprocess
        for _ in range(5):
            deep.gather(torch.zeros(8, dtype=torch.int64, device=dev), torch.bfloat16)
        ids0 = batch()[0]
        torch.cuda.synchronize()
        ts = []
        for _ in range(10):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            deep.gather(ids0, torch.bfloat16)
            b.record()
            torch.cuda.synchronize()
            ts.append(round(a.elapsed_time(b) * 1e3, 1))
        print("standalone deep gather us:", ts, flush=True)
        print(subprocess.run(["nvidia-smi", "--query-gpu=index,pstate,clocks.sm,clocks.mem,power.draw,"
                              "utilization.gpu", "--format=csv"], capture_output=True, text=True).stdout,
              flush=True)
        print("CUDA_VISIBLE_DEVICES", os.environ.get("CUDA_VISIBLE_DEVICES"), "ptr",
              hex(deep.table.weight[0]), flush=True)
        from torch.profiler import ProfilerActivity, profile
        with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
            for _ in range(5):
                step()
            torch.cuda.synchronize()
        print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=25, max_name_column_width=70),
              flush=True)
        print(prof.key_averages().table(sort_by="cpu_time_total", row_limit=12, max_name_column_width=70),
              flush=True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.steps):
        loss = step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / args.steps
    out = {"role": "worker", "index": rank_w, "ms_per_step": round(ms, 3),
           "samples_per_s_this_worker": round(B / ms * 1e3, 1), "workers": n_workers, "ps": n_ps,
           "ids_per_step": B * args.fields, "loss": round(float(loss), 4),
           "kernel_launches_per_step": dp.kernel_launches_per_step}
    print(json.dumps(out), flush=True)
    if not local:
        torch.cuda.synchronize()
        rdv.barrier(prefix + "/done", n_workers)      # nobody leaves while a peer still pushes
        if comm_w is not None:
            comm_w.close()
    return 0


if __name__ == "__main__":
    sys.exit(main())
