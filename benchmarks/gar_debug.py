#!/usr/bin/env python
"""Debug driver for the fused GEMM+all-reduce kernel on real ranks: short spin time-out, prints
after every step, dumps the counters."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tf_operator_b200.ops import _lib  # noqa: E402
from tf_operator_b200.runtime.symm import SymmetricComm  # noqa: E402

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank)
dist.init_process_group("nccl", device_id=torch.device("cuda", rank))
_lib.load().tfop_gemm_allreduce_set_timeout(4_000_000_000)
for (m, n, k) in [(128, 256, 64), (384, 520, 256), (4096, 1024, 8192)]:
    ws = SymmetricComm.gemm_allreduce_workspace_bytes(m, n, world)
    comm = SymmetricComm.from_torch_distributed(ws + m * n * 2 + 4096)
    torch.manual_seed(rank)
    a = (torch.randn(m, k, device="cuda") * 0.1).bfloat16()
    b = (torch.randn(n, k, device="cuda") * 0.1).bfloat16()
    ref = a.float() @ b.float().t()
    dist.all_reduce(ref)
    torch.cuda.synchronize()
    print(f"[{rank}] shape {m}x{n}x{k}: launching", flush=True)
    for it in range(3):
        out = comm.gemm_all_reduce(a, b, ws_off=0, out_off=ws, scale=1.0)
        torch.cuda.synchronize()
        err = (out.float() - ref).abs().max().item()
        tiles = ((m + 127) // 128) * ((n + 255) // 256)
        tpo = (tiles + world - 1) // world
        cnt = comm.view(world * tpo * 128 * 256 * 2, tpo + 2, torch.int32).cpu().tolist()
        print(f"[{rank}] it {it} err {err:.4f} timeout_flag {comm.check_error()} "
              f"arrived {cnt[:min(4, tpo)]} delivered {cnt[tpo]} gen {cnt[tpo + 1]}", flush=True)
    dist.barrier()
    comm.close()
dist.destroy_process_group()
