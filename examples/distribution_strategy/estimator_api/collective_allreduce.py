#!/usr/bin/env python
"""CPU collective all-reduce training — the reference's Estimator + CollectiveAllReduceStrategy
example (examples/v1/distribution_strategy/estimator-API/keras_model_to_estimator.py:26-73):
TFJob 3 Workers, Dense(10->16)->Dense(16->1) = 193 parameters, batch 32, CPU ring all-reduce
(gloo here, gRPC ring in the reference), evaluation on the chief at the end."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))))
from tf_operator_b200.runtime.env import ReplicaContext  # noqa: E402


def main():
    ctx = ReplicaContext.from_env()
    if ctx.world_size > 1:
        ctx.init_process_group("gloo")
    torch.manual_seed(0)
    model = torch.nn.Sequential(torch.nn.Linear(10, 16), torch.nn.ReLU(), torch.nn.Linear(16, 1))
    assert sum(p.numel() for p in model.parameters()) == 193
    opt = torch.optim.SGD(model.parameters(), lr=0.05)
    g = torch.Generator().manual_seed(10 + ctx.rank)
    w = torch.randn(10, 1, generator=torch.Generator().manual_seed(3))
    for step in range(200):
        x = torch.randn(32, 10, generator=g)
        y = x @ w
        opt.zero_grad()
        loss = torch.nn.functional.mse_loss(model(x), y)
        loss.backward()
        if ctx.world_size > 1:
            for p in model.parameters():
                dist.all_reduce(p.grad)
                p.grad /= ctx.world_size
        opt.step()
    x = torch.randn(256, 10, generator=torch.Generator().manual_seed(99))
    ev = torch.nn.functional.mse_loss(model(x), x @ w).item()
    print(f"task {ctx.task_type}:{ctx.task_index} rank {ctx.rank}/{ctx.world_size} eval loss {ev:.4f}", flush=True)
    return 0 if ev < 1.0 else 3


if __name__ == "__main__":
    sys.exit(main())
