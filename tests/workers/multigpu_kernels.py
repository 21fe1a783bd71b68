#!/usr/bin/env python
"""Replica body of tests/test_multigpu_gpu.py::test_collective_kernels: every multi-rank kernel of
the library on REAL peer mappings (one process per GPU, handles through the launcher's rendezvous
store, multicast descriptor through its broker), each checked against an fp32 sum of the ranks'
inputs and for bitwise equality across ranks. Writes <out>/rank<r>.json."""
import hashlib
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from tf_operator_b200.runtime.env import ReplicaContext  # noqa: E402
from tf_operator_b200.runtime.symm import ONE_SHOT, TWO_SHOT, SymmetricComm  # noqa: E402


def digest(t: torch.Tensor) -> str:
    return hashlib.sha1(t.detach().contiguous().view(torch.uint8).cpu().numpy().tobytes()).hexdigest()


def rank_input(rank, n, seed, dtype):
    g = torch.Generator(device="cuda").manual_seed(1000 * seed + rank)
    return torch.randn(n, device="cuda", generator=g).to(dtype)


def main():
    out_dir = sys.argv[1]
    ctx = ReplicaContext.from_env()
    torch.cuda.set_device(0)
    W, R = ctx.world_size, ctx.rank
    comm = ctx.symmetric_comm(64 << 20, nvls_bytes=32 << 20)
    res = {"rank": R, "world": W, "nvls": comm.nvls_bytes > 0, "generation": comm.generation, "checks": {}}

    def ref_sum(n, seed, dtype, scale):
        acc = torch.zeros(n, device="cuda", dtype=torch.float32)
        for r in range(W):
            acc += rank_input(r, n, seed, dtype).float()
        return acc * scale

    def record(name, got, want, atol, rtol):
        torch.cuda.synchronize()
        err = (got.float() - want).abs().max().item()
        tol = atol + rtol * want.abs().max().item()
        res["checks"][name] = {"max_err": err, "tol": tol, "ok": bool(err <= tol), "sha1": digest(got),
                               "error_word_after": bool(comm.check_error())}

    seed = 0
    # ---- P2P one-shot / two-shot, bf16 and fp32 wire, out of place --------------------------
    for algo, aname in ((ONE_SHOT, "oneshot"), (TWO_SHOT, "twoshot")):
        for wire in (torch.bfloat16, torch.float32):
            for n in (1000, 256 * 1024 + 24):
                seed += 1
                x = rank_input(R, n, seed, wire)
                y = torch.empty(n, device="cuda", dtype=torch.float32)
                comm.all_reduce(n, inp=x, out=y, wire_dtype=wire, scale=0.5, algo=algo)
                record(f"p2p_{aname}_{str(wire).split('.')[-1]}_{n}", y, ref_sum(n, seed, wire, 0.5),
                       1e-5, 2e-2 if wire == torch.bfloat16 else 1e-5)
    # ---- two-shot in place on the symmetric buffer (the DDP bucket mode) ---------------------
    n = 1 << 20
    seed += 1
    buf = comm.view(0, n, torch.bfloat16)
    buf.copy_(rank_input(R, n, seed, torch.bfloat16))
    torch.cuda.synchronize()
    comm._exit_barrier()                       # every rank's bucket is filled before anyone reduces
    comm.all_reduce(n, sym_off=0, wire_dtype=torch.bfloat16, scale=1.0 / W, algo=TWO_SHOT)
    record("p2p_twoshot_inplace_bf16", buf, ref_sum(n, seed, torch.bfloat16, 1.0 / W), 1e-5, 2e-2)
    # ---- NVLS: out of place and IN PLACE (what the bench's gradient bucket uses) -------------
    if comm.nvls_bytes:
        for wire in (torch.bfloat16, torch.float32):
            seed += 1
            n = 300 * 1024 + 8
            x = rank_input(R, n, seed, wire)
            y = torch.empty(n, device="cuda", dtype=torch.float32)
            comm.all_reduce_nvls(n, inp=x, out=y, wire_dtype=wire, scale=0.25)
            record(f"nvls_outofplace_{str(wire).split('.')[-1]}", y, ref_sum(n, seed, wire, 0.25),
                   1e-5, 2e-2 if wire == torch.bfloat16 else 1e-5)
        seed += 1
        n = 4 << 20
        nv = comm.nvls_view(n, torch.bfloat16)
        nv.copy_(rank_input(R, n, seed, torch.bfloat16))
        torch.cuda.synchronize()
        comm._exit_barrier()
        comm.all_reduce_nvls(n, sym_off=0, wire_dtype=torch.bfloat16, scale=1.0 / W)
        record("nvls_inplace_bf16", nv, ref_sum(n, seed, torch.bfloat16, 1.0 / W), 1e-5, 2e-2)
        # second call on the reduced data: generation counters advance, no stale flags
        torch.cuda.synchronize()
        comm._exit_barrier()
        want2 = nv.float().clone() * W * 0.5
        torch.cuda.synchronize()
        comm._exit_barrier()
        comm.all_reduce_nvls(n, sym_off=0, wire_dtype=torch.bfloat16, scale=0.5)
        record("nvls_inplace_bf16_second_call", nv, want2, 1e-5, 2e-2)
    # ---- fused GEMM + all-reduce, both flavours ------------------------------------------------
    m, nn, k = 512, 256, 1024
    seed += 1
    a = rank_input(R, m * k, seed, torch.bfloat16).view(m, k) * 0.1
    b = rank_input(R, nn * k, seed + 100, torch.bfloat16).view(nn, k) * 0.1
    want = torch.zeros(m, nn, device="cuda")
    for r in range(W):
        ar = rank_input(r, m * k, seed, torch.bfloat16).view(m, k) * 0.1
        br = rank_input(r, nn * k, seed + 100, torch.bfloat16).view(nn, k) * 0.1
        want += ar.float() @ br.float().t()
    ws = SymmetricComm.gemm_allreduce_workspace_bytes(m, nn, W)
    out_off = (ws + 255) // 256 * 256

    def fresh_workspace():
        # the fused kernels' workspace (partial-tile slots + monotone arrival counters) must start
        # at zero and serve ONE (M, N, world) and one flavour; the collectives
        # above used the same bytes as staging. Every rank zeroes its own, then all meet.
        torch.cuda.synchronize()
        comm._exit_barrier()
        comm.view(0, out_off, torch.uint8).zero_()
        torch.cuda.synchronize()
        comm._exit_barrier()

    fresh_workspace()
    got = comm.gemm_all_reduce(a, b, ws_off=0, out_off=out_off)
    record("gemm_allreduce_p2p", got, want, 2e-2, 2e-2)
    if comm.nvls_bytes:
        from tf_operator_b200.ops import _lib
        lib = _lib.load()
        # the cta_group::2 pipeline with reducer warps, both tile widths, twice on one set of
        # counters (generation 1 and 2), then the weight-gradient operand layout (both MN-major),
        # then the first-generation kernel
        for tile_n in (128, 256):
            lib.tfop_gemm_allreduce_set_tile_n(tile_n)
            fresh_workspace()
            for call in (1, 2):
                got = comm.gemm_all_reduce_nvls(a, b, ws_off=0, out_off=0)
                record(f"gemm_allreduce_nvls_tile{tile_n}_call{call}", got, want, 2e-2, 2e-2)
                torch.cuda.synchronize()
                comm._exit_barrier()
        lib.tfop_gemm_allreduce_set_tile_n(0)
        fresh_workspace()
        got = comm.gemm_all_reduce_nvls(a.t().contiguous(), b.t().contiguous(), ws_off=0, out_off=0,
                                        a_mn=True, b_mn=True)
        record("gemm_allreduce_nvls_wgrad_layout", got, want, 2e-2, 2e-2)
        torch.cuda.synchronize()
        comm._exit_barrier()
        lib.tfop_gemm_allreduce_set_legacy(1)
        fresh_workspace()
        got = comm.gemm_all_reduce_nvls(a, b, ws_off=0, out_off=0)
        record("gemm_allreduce_nvls", got, want, 2e-2, 2e-2)
        lib.tfop_gemm_allreduce_set_legacy(0)
    res["error_word"] = comm.check_error()
    with open(os.path.join(out_dir, f"rank{R}.json"), "w") as f:
        json.dump(res, f)
    assert comm.close()
    bad = {k: v for k, v in res["checks"].items() if not v["ok"]}
    print(f"rank {R}: {len(res['checks']) - len(bad)}/{len(res['checks'])} checks ok; "
          f"error_word={res['error_word']}; failed: {json.dumps(bad)}", flush=True)


if __name__ == "__main__":
    main()
