#!/usr/bin/env python
"""Replica body of tests/test_multigpu_gpu.py::test_fused_wgrad_*: three AdamW steps of a small
TcLinear stack through FlatDataParallel over the NVLS gradient buffer, once with the weight
gradients summed by the bucket all-reduce and once by the fused GEMM -> in-switch reduction
(argv[2] = bucket | fused). Dumps the master weights by parameter name."""
import hashlib
import json
import os
import sys

import torch
import torch.nn as nn

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from tf_operator_b200.ops.gemm import TcLinear  # noqa: E402
from tf_operator_b200.parallel.engine import FlatDataParallel  # noqa: E402
from tf_operator_b200.runtime.env import ReplicaContext  # noqa: E402


def main():
    out_dir, mode = sys.argv[1], sys.argv[2]
    ctx = ReplicaContext.from_env()
    torch.cuda.set_device(0)
    torch.manual_seed(1234)
    model = nn.Sequential(TcLinear(512, 1024, act="gelu"), TcLinear(1024, 768, act="relu"),
                          TcLinear(768, 384), nn.LayerNorm(384), TcLinear(384, 16)).to(torch.bfloat16)
    nparam = sum((p.numel() + 7) // 8 * 8 for p in model.parameters())
    comm = ctx.symmetric_comm(nparam * 2 + (8 << 20), nvls_bytes=nparam * 2 + (1 << 20))
    assert comm.nvls_bytes > 0, "this test needs NVLS"
    eng = FlatDataParallel(model, comm=comm, optimizer="adamw", lr=1e-3, weight_decay=0.01,
                           use_nvls=True, fused_wgrad=(mode == "fused"), bucket_bytes=1 << 20)
    assert eng.fused_wgrad == (mode == "fused")
    g = torch.Generator(device="cuda").manual_seed(5 + ctx.rank)
    x = torch.randn(1024, 512, device="cuda", generator=g).bfloat16()
    y = torch.randn(1024, 16, device="cuda", generator=g)
    losses = [float(eng.train_step(lambda: ((model(x).float() - y) ** 2).mean())) for _ in range(3)]
    torch.cuda.synchronize()
    comm.raise_if_failed()
    w = {s.name: eng.flat.master[s.offset: s.offset + s.numel].detach().float().cpu()
         for s in eng.flat.slots}
    torch.save(w, os.path.join(out_dir, f"{mode}_rank{ctx.rank}.pt"))
    flat = torch.cat([w[k] for k in sorted(w)])
    with open(os.path.join(out_dir, f"{mode}_rank{ctx.rank}.json"), "w") as f:
        json.dump({"losses": losses, "fused_params": sum(1 for s in eng.flat.slots if s.offset < eng.flat.n_fused),
                   "buckets": len(eng.flat.buckets),
                   "sha1": hashlib.sha1(flat.numpy().tobytes()).hexdigest()}, f)
    assert comm.close()
    print(f"{mode} rank {ctx.rank}: losses {losses}", flush=True)


if __name__ == "__main__":
    main()
