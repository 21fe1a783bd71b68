#!/usr/bin/env python
"""Replica body of tests/test_multigpu_gpu.py::test_ddp_*: W ranks x batch B, three SGD steps of
ResNet-18 through DataParallelEngine (NVLS or P2P gradient bucket), weights dumped per rank.
With WORLD_SIZE unset it runs the single-process twin on the concatenated batch W*B."""
import hashlib
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from tf_operator_b200.models.resnet import ResNet  # noqa: E402
from tf_operator_b200.ops import preproc  # noqa: E402
from tf_operator_b200.parallel.engine import DataParallelEngine, FlatDataParallel  # noqa: E402
from tf_operator_b200.runtime.env import ReplicaContext  # noqa: E402

B, S, STEPS = 8, 64, 3


def batch(rank):
    g = torch.Generator().manual_seed(77 + rank)
    x = torch.randint(0, 256, (B, S, S, 3), dtype=torch.uint8, generator=g)
    y = torch.randint(0, 10, (B,), dtype=torch.int64, generator=g)
    return x, y


def main():
    out_dir, world_twin, mode = sys.argv[1], int(sys.argv[2]), sys.argv[3]
    ctx = ReplicaContext.from_env()
    torch.cuda.set_device(0)
    torch.manual_seed(1234)
    model = ResNet(18, num_classes=10, width=64)
    comm = None
    if ctx.world_size > 1:
        nparam = sum((p.numel() + 7) // 8 * 8 for p in model.parameters())
        comm = ctx.symmetric_comm(nparam * 2 + (8 << 20),
                                  nvls_bytes=nparam * 2 + (1 << 20) if mode == "nvls" else 0)
    use_nvls = comm is not None and comm.nvls_bytes > 0
    if ctx.world_size > 1:
        xs, ys = batch(ctx.rank)
        eng = DataParallelEngine(model, comm=comm, batch_size=B, image_size=S, lr=0.05, graph=False,
                                 use_nvls=use_nvls, bucket_bytes=4 << 20)
        xs, ys = xs.pin_memory(), ys.pin_memory()
        losses = [eng.step(xs, ys).result() for _ in range(STEPS)]
    else:
        # The twin: one process, the same W micro-batches per step, each with its own BatchNorm
        # statistics (as on W replicas), losses averaged, ONE backward and one optimizer step of
        # the same fused SGD kernel - the definition of what data parallelism must reproduce.
        eng = FlatDataParallel(model, optimizer="sgd", lr=0.05, momentum=0.9, weight_decay=1e-4,
                               channels_last=True)
        parts = [tuple(t.cuda() for t in batch(r)) for r in range(world_twin)]
        bufs = [torch.zeros((B, S, S, 3), dtype=torch.bfloat16, device="cuda") for _ in parts]

        def closure():
            total = 0.0
            for (xu8, y), buf in zip(parts, bufs):
                x = preproc.u8_to_bf16_normalize(xu8, buf)
                total = total + torch.nn.functional.cross_entropy(model(x).float(), y)
            return total / world_twin
        losses = [float(eng.train_step(closure)) for _ in range(STEPS)]
    torch.cuda.synchronize()
    w = eng.flat.master.detach().float().cpu()
    name = f"rank{ctx.rank}" if ctx.world_size > 1 else "twin"
    torch.save(w, os.path.join(out_dir, f"{name}.pt"))
    with open(os.path.join(out_dir, f"{name}.json"), "w") as f:
        json.dump({"losses": losses, "nvls": use_nvls, "world": ctx.world_size,
                   "sha1": hashlib.sha1(w.numpy().tobytes()).hexdigest()}, f)
    if comm is not None:
        assert comm.close()
    print(f"{name}: losses {losses}", flush=True)


if __name__ == "__main__":
    main()
