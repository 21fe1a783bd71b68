#!/usr/bin/env python
"""Fused GEMM -> reduce-scatter -> all-gather (one kernel over peer memory) vs the unfused
pipelines: tcgen05 GEMM + fused P2P/NVLS all-reduce, and cuBLAS GEMM + NCCL all-reduce.

    python -m torch.distributed.run --nproc-per-node N ... benchmarks/gemm_allreduce_bench.py

Roofline per shape: slower of (2*M*N*K / measured GEMM peak) and (bytes that must cross
NVLink per GPU = 2*(W-1)/W * M*N*2 / 770 GB/s).
"""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from tf_operator_b200.ops import _lib  # noqa: E402
from tf_operator_b200.ops.gemm import gemm_tn  # noqa: E402
from tf_operator_b200.runtime.symm import TWO_SHOT, SymmetricComm  # noqa: E402


def timed(fn, iters=20, warm=5):
    for _ in range(warm):
        fn()
    dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters], device="cuda", dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item()) * 1e3


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    peaks = {"bf16_tflops": 1630.6}
    pk = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
    if os.path.exists(pk):
        peaks = json.load(open(pk))
    shapes = [(4096, 1024, 8192), (1024, 4096, 8192), (1024, 1024, 8192), (3072, 1024, 8192),
              (4096, 4096, 4096), (8192, 8192, 2048)]
    max_out = max(m * n for m, n, _ in shapes) * 2
    max_ws = max(SymmetricComm.gemm_allreduce_workspace_bytes(m, n, world) for m, n, _ in shapes)
    ws_off, out_off = 0, (max_ws + 255) // 256 * 256
    comm = SymmetricComm.from_torch_distributed(out_off + max_out + (1 << 20))
    nvls = comm.enable_nvls(max_out + (1 << 20))
    res = {"world": world, "nvls": bool(nvls), "rows": []}

    def say(msg):
        if rank == 0:
            print(msg, file=sys.stderr, flush=True)
    say(f"comm ready, nvls={nvls}")
    ok = True
    for (m, n, k) in shapes:
        # the workspace's counters are tied to one (M, N): zero it for every new shape
        torch.cuda.synchronize()
        dist.barrier()
        comm.view(ws_off, out_off - ws_off, torch.uint8).zero_()
        torch.cuda.synchronize()
        dist.barrier()
        torch.manual_seed(10 + rank)
        a = (torch.randn(m, k, device="cuda") * 0.1).bfloat16()
        b = (torch.randn(n, k, device="cuda") * 0.1).bfloat16()
        ref = a.float() @ b.float().t()
        dist.all_reduce(ref)
        ref /= world
        out = comm.gemm_all_reduce(a, b, ws_off=ws_off, out_off=out_off, scale=1.0 / world)
        torch.cuda.synchronize()
        dist.barrier()
        err = (out.float() - ref).abs().max().item()
        tol = 0.02 * ref.abs().max().item() + 0.05
        good = err < tol and not comm.check_error()
        ok &= good
        row = {"m": m, "n": n, "k": k, "max_err": err, "ok": bool(good)}
        say(f"{m}x{n}x{k}: fused correct={good} err={err:.4f}")
        row["fused_us"] = round(timed(lambda: comm.gemm_all_reduce(
            a, b, ws_off=ws_off, out_off=out_off, scale=1.0 / world)), 1)
        say(f"  fused {row['fused_us']} us")
        if nvls:
            torch.cuda.synchronize()
            dist.barrier()
            comm.view(ws_off, out_off - ws_off, torch.uint8).zero_()
            torch.cuda.synchronize()
            dist.barrier()
            outn = comm.gemm_all_reduce_nvls(a, b, ws_off=ws_off, scale=1.0 / world)
            torch.cuda.synchronize()
            dist.barrier()
            errn = (outn.float() - ref).abs().max().item()
            goodn = errn < tol and not comm.check_error()
            ok &= goodn
            row["fused_nvls_max_err"] = errn
            # the counters of the two flavours share ws_off: zero them before switching back
            row["fused_nvls_us"] = round(timed(lambda: comm.gemm_all_reduce_nvls(
                a, b, ws_off=ws_off, scale=1.0 / world)), 1)
            say(f"  fused nvls correct={goodn} err={errn:.4f} {row['fused_nvls_us']} us")
            # the same kernel at both tile widths, and the first-generation kernel (one CTA per
            # tile, 3 stages); the counters' layout follows the tile count: zero in between
            lib = _lib.load()
            for name, tile_n, legacy in (("fused_nvls_tile128_us", 128, 0), ("fused_nvls_tile256_us", 256, 0),
                                         ("fused_nvls_first_generation_us", 0, 1)):
                lib.tfop_gemm_allreduce_set_tile_n(tile_n)
                lib.tfop_gemm_allreduce_set_legacy(legacy)
                torch.cuda.synchronize()
                dist.barrier()
                comm.view(ws_off, out_off - ws_off, torch.uint8).zero_()
                torch.cuda.synchronize()
                dist.barrier()
                row[name] = round(timed(lambda: comm.gemm_all_reduce_nvls(
                    a, b, ws_off=ws_off, scale=1.0 / world)), 1)
                ok &= not comm.check_error()
            lib.tfop_gemm_allreduce_set_tile_n(0)
            lib.tfop_gemm_allreduce_set_legacy(0)
            torch.cuda.synchronize()
            dist.barrier()
            comm.view(ws_off, out_off - ws_off, torch.uint8).zero_()
            torch.cuda.synchronize()
            dist.barrier()
            say(f"  tile128 {row['fused_nvls_tile128_us']} tile256 {row['fused_nvls_tile256_us']} "
                f"first-generation {row['fused_nvls_first_generation_us']} us")
        # unfused, all ours: tcgen05 GEMM into symmetric memory, then in-place fused all-reduce
        sym_out = comm.view(out_off, m * n, torch.bfloat16).view(m, n)

        def ours_unfused():
            gemm_tn(a, b, out=sym_out)
            comm.all_reduce(m * n, sym_off=out_off, scale=1.0 / world, algo=TWO_SHOT)
        row["tcgen05_gemm_plus_p2p_allreduce_us"] = round(timed(ours_unfused), 1)
        say(f"  unfused p2p {row['tcgen05_gemm_plus_p2p_allreduce_us']} us")
        if nvls:
            nv = comm.nvls_view(m * n, torch.bfloat16).view(m, n)

            def ours_nvls():
                gemm_tn(a, b, out=nv)
                comm.all_reduce_nvls(m * n, scale=1.0 / world)
            row["tcgen05_gemm_plus_nvls_allreduce_us"] = round(timed(ours_nvls), 1)
            say(f"  unfused nvls {row['tcgen05_gemm_plus_nvls_allreduce_us']} us")
        c = torch.empty(m, n, device="cuda", dtype=torch.bfloat16)

        def stock():
            torch.matmul(a, b.t(), out=c)
            dist.all_reduce(c)
        row["cublas_gemm_plus_nccl_us"] = round(timed(stock), 1)
        say(f"  cublas+nccl {row['cublas_gemm_plus_nccl_us']} us")
        row["gemm_only_tcgen05_us"] = round(timed(lambda: gemm_tn(a, b, out=sym_out)), 1)
        t_gemm = 2.0 * m * n * k / (peaks["bf16_tflops"] * 1e12) * 1e6
        t_link = 2.0 * (world - 1) / world * m * n * 2 / 770e9 * 1e6
        row["roofline_us"] = round(max(t_gemm, t_link), 1)
        best = min(row["fused_us"], row.get("fused_nvls_us", 1e9))
        row["fused_frac_of_roofline"] = round(row["roofline_us"] / best, 3)
        res["rows"].append(row)
    res["ok"] = bool(ok)
    if rank == 0:
        print(json.dumps(res))
        if args.out:
            json.dump(res, open(args.out, "w"), indent=1)
    dist.barrier()
    comm.close()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
