#!/usr/bin/env python
"""ResNet-50 data-parallel training as a PyTorchJob replica (BASELINE.json config 2:
1 Master + 7 Workers, one B200 per replica, bf16, synthetic 224x224 images).

The replica reads the operator's env contract (MASTER_ADDR / MASTER_PORT / WORLD_SIZE / RANK,
pkg/controller.v1/pytorch/pytorch.go:41-65), maps its peers' symmetric gradient buffers through
the launcher's rendezvous store and trains with the fused P2P all-reduce — NCCL is only used
with ``--impl nccl`` (the measured baseline).
"""
import argparse
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))

from tf_operator_b200.models.resnet import ResNet  # noqa: E402
from tf_operator_b200.parallel.engine import DataParallelEngine  # noqa: E402
from tf_operator_b200.runtime.env import ReplicaContext  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--batch-size", type=int, default=256)
    ap.add_argument("--image-size", type=int, default=224)
    ap.add_argument("--depth", type=int, default=50)
    ap.add_argument("--lr", type=float, default=0.1)
    args = ap.parse_args()

    ctx = ReplicaContext.from_env()
    torch.cuda.set_device(0)            # the launcher exposes exactly one GPU per replica
    torch.backends.cudnn.benchmark = True
    torch.manual_seed(1234)             # identical initial weights on every replica
    model = ResNet(args.depth)
    nparam = sum((p.numel() + 7) // 8 * 8 for p in model.parameters())
    comm = ctx.symmetric_comm(nparam * 2 + (8 << 20)) if ctx.world_size > 1 else None
    eng = DataParallelEngine(model, comm=comm, batch_size=args.batch_size, image_size=args.image_size,
                             lr=args.lr * ctx.world_size, stem_s2d=True)
    torch.manual_seed(100 + ctx.rank)
    B, S = args.batch_size, args.image_size
    xs = [torch.randint(0, 256, (B, S, S, 3), dtype=torch.uint8).pin_memory() for _ in range(2)]
    ys = [torch.randint(0, 1000, (B,), dtype=torch.int64).pin_memory() for _ in range(2)]
    print(f"rank {ctx.rank}/{ctx.world_size} on {torch.cuda.get_device_name(0)} "
          f"master={ctx.master_addr}:{ctx.master_port}", flush=True)
    for _ in range(5):
        eng.step(xs[0], ys[0]).result()
    torch.cuda.synchronize()
    t0 = time.time()
    pending = None
    for i in range(args.steps):
        h = eng.step(xs[i % 2], ys[i % 2])
        if pending is not None:
            loss = pending.result()
            if ctx.rank == 0 and i % 10 == 0:
                print(f"step {i} loss {loss:.4f}", flush=True)
        pending = h
    loss = pending.result()
    dt = time.time() - t0
    if ctx.rank == 0:
        print(json.dumps({"img_per_s": round(args.steps * B * ctx.world_size / dt, 1),
                          "ms_per_step": round(dt / args.steps * 1e3, 3), "world": ctx.world_size,
                          "final_loss": round(loss, 4)}), flush=True)
    if comm is not None:
        torch.cuda.synchronize()
        time.sleep(0.5)                 # peers may still be reading this rank's buffers
    return 0


if __name__ == "__main__":
    sys.exit(main())
