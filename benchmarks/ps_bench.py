#!/usr/bin/env python
"""Parameter-server pull / push, piece by piece, with the table LOCAL and with it in a PEER's HBM
over NVLink (BASELINE.json config 4 shapes: 8192 x 26 ids per step out of 10M rows of dim 64).

    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 benchmarks/ps_bench.py

Rank 0 owns the shard and sleeps on the host (its GPU runs nothing, like a PS replica); rank 1
is the worker and times (CUDA events, 20 calls after 5):
  gather                       K4, rows out of the peer's HBM with P2P loads
  push, device-side merge      K5b with plain load + two RED (default) / with fetch-and-add
  push, library sort + K5      torch.sort + unique_consecutive + cumsum, then one kernel
and the same against a local table. Stock comparator for the pull+push pair: NCCL send/recv of ids,
rows and gradients plus torch.unique / index_add_ on the owner (what wide_deep_baseline.py does).
"""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tf_operator_b200.ops import _lib  # noqa: E402
from tf_operator_b200.ops.ps_embedding import ShardedEmbedding  # noqa: E402
from tf_operator_b200.runtime.symm import SymmetricComm  # noqa: E402


def timed(fn, iters=20, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return round(e0.elapsed_time(e1) / iters * 1e3, 1)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=10_000_000)
    ap.add_argument("--dim", type=int, default=64)
    ap.add_argument("--ids", type=int, default=8192 * 26)
    ap.add_argument("--out", default="")
    ap.add_argument("--no-nccl-first", action="store_true",
                    help="exchange handles over gloo and measure before NCCL is initialised")
    ap.add_argument("--restrict", action="store_true",
                    help="each rank sees only its own GPU (CUDA_VISIBLE_DEVICES), like a launched replica")
    args = ap.parse_args()
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    assert world == 2
    if args.restrict:
        # what the launcher does for a replica with nvidia.com/gpu: 1 - the process sees one GPU
        os.environ["CUDA_VISIBLE_DEVICES"] = os.environ.get("LOCAL_RANK", str(rank))
        torch.cuda.set_device(0)
    else:
        torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
    dev = torch.device("cuda", torch.cuda.current_device())
    if args.no_nccl_first:
        # the peer mapping is made and measured BEFORE any NCCL communicator exists between the
        # two processes (a launched replica never creates one)
        dist.init_process_group("gloo")
    else:
        dist.init_process_group("nccl", device_id=dev)
    nbytes = ShardedEmbedding.shard_bytes(args.rows, args.dim)
    comm = SymmetricComm.from_torch_distributed(nbytes if rank == 0 else (1 << 20))
    if rank == 0:
        ShardedEmbedding.init_local_shard(comm, args.rows, args.dim, 0, seed=0)
    torch.cuda.synchronize()
    dist.barrier()
    res = {"rows": args.rows, "dim": args.dim, "ids": args.ids, "one_gpu_visible_per_rank": args.restrict}
    lib = _lib.load()
    host = dist.new_group(backend="gloo")          # waits that leave the GPUs alone
    if rank == 1:
        g = torch.Generator(device=dev).manual_seed(3)
        ids = (torch.rand(args.ids, device=dev, generator=g).pow(3.0) * (args.rows - 1)).long()
        grad = torch.randn(args.ids, args.dim, device=dev, generator=g) * 1e-3
        res["unique_ids"] = int(torch.unique(ids).numel())
        remote = ShardedEmbedding.from_comm(comm, [0], args.rows, args.dim, 0, lr=0.05)
        local = ShardedEmbedding.local(args.rows, args.dim, 1, dev, lr=0.05)

    def measure(emb):
        r = {}
        r["gather_us"] = timed(lambda: emb.gather(ids, torch.bfloat16))
        lib.tfop_ps_set_push_mode(0)
        r["push_merge_on_device_load_red_us"] = timed(lambda: emb.push(ids, grad))
        lib.tfop_ps_set_push_mode(1)
        r["push_merge_on_device_fetch_add_us"] = timed(lambda: emb.push(ids, grad))
        lib.tfop_ps_set_push_mode(0)
        r["push_library_sort_us"] = timed(lambda: emb.push_sorted(ids, grad))
        row_bytes = args.dim * 4
        r["gather_GBps"] = round(args.ids * row_bytes / r["gather_us"] / 1e3, 1)
        return r

    # phase A: the owner's GPU has NOTHING to run (its process sleeps in a host-side barrier)
    torch.cuda.synchronize()
    dist.barrier(group=host)
    if rank == 1:
        import time
        time.sleep(1.0)                              # let the owner's GPU settle into idle
        res["remote_owner_gpu_idle"] = measure(remote)
    dist.barrier(group=host)
    # phase B: local table, then the idle-gap sweep on the remote one
    if rank == 0:
        dist.barrier(group=host)
        dist.barrier(group=host)
    else:
        dist.barrier(group=host)
        import time
        res["local"] = measure(local)
        # the same gather with the GPU and the links left idle for a while before each call, and
        # with ids that change from call to call (what a training step does)
        sweep = {}
        for gap_ms in (0, 1, 3, 10, 50):
            ts = []
            for i in range(12):
                idi = (torch.rand(args.ids, device=dev, generator=g).pow(3.0) * (args.rows - 1)).long()
                torch.cuda.synchronize()
                time.sleep(gap_ms / 1e3)
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record()
                remote.gather(idi, torch.bfloat16)
                b.record()
                torch.cuda.synchronize()
                ts.append(a.elapsed_time(b) * 1e3)
            ts = sorted(ts[2:])
            sweep[f"idle_{gap_ms}ms_before_each_call_us"] = {"median": round(ts[len(ts) // 2], 1),
                                                            "min": round(ts[0], 1), "max": round(ts[-1], 1)}
        res["remote_gather_fresh_ids_vs_idle_gap"] = sweep
        dist.barrier(group=host)
    # ---- stock pull + push over NCCL (ids there, rows back, gradients there, update) -----------
    n = args.ids
    nccl = None
    if args.no_nccl_first:
        nccl = dist.new_group(backend="nccl")
        dist.barrier(group=nccl)
        torch.cuda.synchronize()
        if rank == 1:
            res["remote_after_nccl_init"] = measure(remote)
        dist.barrier(group=host)
    if rank == 0:
        weight = comm.view(0, args.rows * args.dim, torch.float32).view(args.rows, args.dim)
        accum = comm.view(args.rows * args.dim * 4, args.rows * args.dim, torch.float32).view(args.rows, args.dim)
        got_ids = torch.empty(n, dtype=torch.int64, device=dev)
        got_g = torch.empty(n, args.dim, device=dev)

        def stock():
            dist.recv(got_ids, 1, group=nccl)
            dist.send(weight[got_ids], 1, group=nccl)
            dist.recv(got_g, 1, group=nccl)
            uniq, inv = torch.unique(got_ids, return_inverse=True)
            gs = torch.zeros(uniq.numel(), args.dim, device=dev).index_add_(0, inv, got_g)
            acc = accum[uniq] + gs * gs
            accum[uniq] = acc
            weight[uniq] -= 0.05 * gs / (acc.sqrt() + 1e-10)
    else:
        rows = torch.empty(n, args.dim, device=dev)

        def stock():
            dist.send(ids, 0, group=nccl)
            dist.recv(rows, 0, group=nccl)
            dist.send(grad, 0, group=nccl)
    t = torch.tensor([timed(stock)], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=nccl)
    res["stock_nccl_pull_plus_push_us"] = round(float(t.item()), 1)
    if rank == 1:
        r = res["remote_owner_gpu_idle"]
        res["ours_remote_pull_plus_push_us"] = round(r["gather_us"] + r["push_merge_on_device_load_red_us"], 1)
        line = json.dumps(res)
        print(line, flush=True)
        if args.out:
            with open(args.out, "w") as f:
                f.write(line + "\n")
    dist.barrier(group=host)
    comm.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
