#!/usr/bin/env python
"""Multi-GPU check + microbenchmark of the fused P2P all-reduce against NCCL.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29511 benchmarks/allreduce_bench.py [--out gpurun_out/allreduce_N.json]

Correctness: every rank's result must equal NCCL's sum within bf16 tolerance and be
bitwise identical across ranks. Timing: CUDA events, max over ranks, L2 flushed
between iterations is unnecessary here because peer data never sits in the local L2
(peer loads bypass it) — stated in the output.
"""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from tf_operator_b200.runtime.symm import ONE_SHOT, TWO_SHOT, SymmetricComm  # noqa: E402


def timed(fn, iters, warm=5):
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
    return float(t.item()) * 1e3  # us


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default="")
    ap.add_argument("--max-mb", type=int, default=256)
    ap.add_argument("--link-bytes", action="store_true",
                    help="also read the NVLink data counters (nvidia-smi nvlink -gt d) around 200 calls "
                         "at 64 MB, and sweep the block count of the in-switch all-reduce")
    args = ap.parse_args()
    rank = int(os.environ["RANK"])
    world = int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    max_bytes = args.max_mb << 20
    comm = SymmetricComm.from_torch_distributed(2 * max_bytes + (1 << 20))
    nvls = comm.enable_nvls(max_bytes)
    results = {"world": world, "nvls": bool(nvls), "rows": [], "note": "CUDA events, max over ranks; peer loads bypass local L2"}
    ok = True
    sizes = [1 << 10, 8 << 10, 64 << 10, 256 << 10, 1 << 20, 4 << 20, 16 << 20, 64 << 20, max_bytes]
    # histogram shape of BASELINE config 5: 256 bins x 500 features x (g,h) fp32
    hist_n = 256 * 500 * 2
    for nbytes in sizes:
        n = nbytes // 2
        torch.manual_seed(100 + rank)
        x = torch.randn(n, device=dev).bfloat16()
        ref = x.clone().float()
        dist.all_reduce(ref)
        ref = ref / world
        row = {"bytes": nbytes}
        for name, algo in (("oneshot", ONE_SHOT), ("twoshot", TWO_SHOT)):
            if algo == ONE_SHOT and nbytes > (8 << 20):
                continue
            out = torch.empty_like(x)
            comm.all_reduce(n, inp=x, out=out, scale=1.0 / world, algo=algo)
            torch.cuda.synchronize()
            err = (out.float() - ref).abs().max().item()
            gathered = [torch.empty_like(out) for _ in range(world)]
            dist.all_gather(gathered, out)
            same = all(torch.equal(g, gathered[0]) for g in gathered)
            good = err < 0.05 and same and not comm.check_error()
            ok &= good
            us = timed(lambda: comm.all_reduce(n, inp=x, out=out, scale=1.0 / world, algo=algo), 20)
            row[name] = {"us": round(us, 2), "max_err": err, "identical": same,
                         "busbw_GBps": round(2 * (world - 1) / world * nbytes / us / 1e3, 1)}
        # in-place two-shot on symmetric memory (the DDP gradient-bucket mode)
        sym = comm.view(0, n, torch.bfloat16)
        sym.copy_(x)
        dist.barrier()
        torch.cuda.synchronize()
        comm.all_reduce(n, sym_off=0, scale=1.0 / world, algo=TWO_SHOT)
        torch.cuda.synchronize()
        err = (sym.float() - ref).abs().max().item()
        ok &= err < 0.05
        us = timed(lambda: comm.all_reduce(n, sym_off=0, scale=1.0 / world, algo=TWO_SHOT), 20)
        row["twoshot_inplace"] = {"us": round(us, 2), "max_err": err,
                                  "busbw_GBps": round(2 * (world - 1) / world * nbytes / us / 1e3, 1)}
        if nvls:
            out = torch.empty_like(x)
            comm.all_reduce_nvls(n, inp=x, out=out, scale=1.0 / world)
            torch.cuda.synchronize()
            err = (out.float() - ref).abs().max().item()
            gathered = [torch.empty_like(out) for _ in range(world)]
            dist.all_gather(gathered, out)
            same = all(torch.equal(g, gathered[0]) for g in gathered)
            ok &= err < 0.05 and same
            us = timed(lambda: comm.all_reduce_nvls(n, inp=x, out=out, scale=1.0 / world), 20)
            row["nvls"] = {"us": round(us, 2), "max_err": err, "identical": same,
                           "busbw_GBps": round(2 * (world - 1) / world * nbytes / us / 1e3, 1)}
            # in place on the multicast-bound buffer: the mode the gradient buckets use. Checked
            # first (against the fp32 mean, and bitwise across ranks), then timed with scale
            # 1/world so that repeated reductions of the buffer stay bounded (the mean of means)
            nv = comm.nvls_view(n, torch.bfloat16)
            nv.copy_(x)
            dist.barrier()
            torch.cuda.synchronize()
            comm.all_reduce_nvls(n, sym_off=0, scale=1.0 / world)
            torch.cuda.synchronize()
            err = (nv.float() - ref).abs().max().item()
            gathered = [torch.empty_like(nv) for _ in range(world)]
            dist.all_gather(gathered, nv.clone())
            same = all(torch.equal(g, gathered[0]) for g in gathered)
            ok &= err < 0.05 and same and not comm.check_error()
            dist.barrier()
            us = timed(lambda: comm.all_reduce_nvls(n, sym_off=0, scale=1.0 / world), 20)
            row["nvls_inplace"] = {"us": round(us, 2), "max_err": err, "identical": same,
                                   "finite_after_timing": bool(torch.isfinite(nv.float()).all()),
                                   "busbw_GBps": round(2 * (world - 1) / world * nbytes / us / 1e3, 1)}
        y = x.clone()
        us = timed(lambda: dist.all_reduce(y), 20)
        row["nccl"] = {"us": round(us, 2),
                       "busbw_GBps": round(2 * (world - 1) / world * nbytes / us / 1e3, 1)}
        results["rows"].append(row)
    # ---- traffic evidence: NVLink data counters of this rank's GPU around 200 calls at 64 MB ----
    if args.link_bytes:
        import re
        import subprocess

        def link_kib():
            try:
                txt = subprocess.run(["nvidia-smi", "nvlink", "-gt", "d", "-i", str(local)],
                                     capture_output=True, text=True, timeout=30).stdout
            except Exception as e:  # noqa: BLE001
                return None, None, repr(e)
            tx = sum(int(v) for v in re.findall(r"Data Tx:\s*(\d+)\s*KiB", txt))
            rx = sum(int(v) for v in re.findall(r"Data Rx:\s*(\d+)\s*KiB", txt))
            return tx, rx, ("" if (tx or rx) else txt[-300:])
        nbytes = 64 << 20
        n = nbytes // 2
        calls = 200
        torch.manual_seed(100 + rank)
        x = torch.randn(n, device=dev).bfloat16()
        sym = comm.view(0, n, torch.bfloat16)
        modes = [("twoshot_inplace", lambda: comm.all_reduce(n, sym_off=0, scale=1.0 / world, algo=TWO_SHOT),
                  2.0 * (world - 1) / world * nbytes)]
        if nvls:
            nv = comm.nvls_view(n, torch.bfloat16)
            nv.copy_(x)
            modes.append(("nvls_inplace", lambda: comm.all_reduce_nvls(n, sym_off=0, scale=1.0 / world),
                          (1.0 + 1.0 / world) * nbytes))
        y = x.clone()
        modes.append(("nccl", lambda: dist.all_reduce(y), None))
        sym.copy_(x)
        link = {"bytes": nbytes, "calls": calls, "gpu": local, "modes": {}}
        for name, fn, algo_bytes in modes:
            torch.cuda.synchronize()
            dist.barrier()
            tx0, rx0, note = link_kib()
            dist.barrier()
            for _ in range(calls):
                fn()
            torch.cuda.synchronize()
            dist.barrier()
            tx1, rx1, _ = link_kib()
            dist.barrier()
            rec = {"algorithmic_bytes_per_call_and_direction": algo_bytes}
            if tx0 is None or note:
                rec["counters"] = f"unavailable: {note}"
            else:
                rec["measured_tx_bytes_per_call"] = round((tx1 - tx0) * 1024 / calls)
                rec["measured_rx_bytes_per_call"] = round((rx1 - rx0) * 1024 / calls)
            link["modes"][name] = rec
        results["nvlink_counters_rank0" if rank == 0 else "_"] = link
        results.pop("_", None)
        # block-count sweep of the in-switch all-reduce at 64 MB (cap: TFOP_MAX_BLOCKS = 64)
        if nvls:
            sweep = {}
            for blocks in (8, 16, 32, 48, 64):
                us = timed(lambda: comm.all_reduce_nvls(n, sym_off=0, scale=1.0 / world, max_blocks=blocks), 20)
                sweep[str(blocks)] = round(us, 2)
            results["nvls_inplace_64MB_us_by_blocks"] = sweep
    # fp32 histogram all-reduce (K6)
    h = torch.randn(hist_n, device=dev)
    href = h.clone()
    dist.all_reduce(href)
    hout = torch.empty_like(h)
    hrow = {"elements": hist_n, "bytes": hist_n * 4}
    for name, algo in (("oneshot", ONE_SHOT), ("twoshot", TWO_SHOT)):
        comm.all_reduce(hist_n, inp=h, out=hout, wire_dtype=torch.float32, algo=algo)
        torch.cuda.synchronize()
        err = (hout - href).abs().max().item()
        ok &= err < 1e-3
        us = timed(lambda: comm.all_reduce(hist_n, inp=h, out=hout, wire_dtype=torch.float32,
                                           algo=algo), 50)
        hrow[name] = {"us": round(us, 2), "max_err": err}
    if nvls:
        comm.all_reduce_nvls(hist_n, inp=h, out=hout, wire_dtype=torch.float32)
        torch.cuda.synchronize()
        err = (hout - href).abs().max().item()
        ok &= err < 1e-3
        us = timed(lambda: comm.all_reduce_nvls(hist_n, inp=h, out=hout, wire_dtype=torch.float32), 50)
        hrow["nvls"] = {"us": round(us, 2), "max_err": err}
    hy = h.clone()
    hrow["nccl"] = {"us": round(timed(lambda: dist.all_reduce(hy), 50), 2)}
    results["histogram_f32"] = hrow
    results["ok"] = bool(ok)
    if rank == 0:
        print(json.dumps(results))
        if args.out:
            with open(args.out, "w") as f:
                json.dump(results, f, indent=1)
    dist.barrier()
    comm.close()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
