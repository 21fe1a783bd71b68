#!/usr/bin/env python
"""Distributed MNIST with parameter servers — the TFJob dist-mnist workload
(BASELINE.json config 1: 1 PS + 2 Workers, CPU plumbing).

Same job shape and training recipe as the reference's examples/v1/dist-mnist/dist_mnist.py
(:97-299): parse TF_CONFIG; PS replicas serve forever; workers hold no variables — every step
they pull the 784->100->10 MLP (79,510 fp32 parameters, vars round-robined over the PS shards),
compute the clipped cross-entropy gradient on a batch of 100, push it, and the PS applies Adam
(lr 0.01), asynchronously, until ``global_step >= train_steps``; then worker 0 reports the
validation cross-entropy. Data is synthetic MNIST-shaped (no dataset download on this box).
"""
import argparse
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))

from tf_operator_b200.runtime.env import ReplicaContext  # noqa: E402
from tf_operator_b200.runtime.ps import ParameterServer, PSClient  # noqa: E402

IMAGE_PIXELS = 28


def synthetic_mnist(n, seed):
    """Linearly separable-ish 10-class data of MNIST shape."""
    g = torch.Generator().manual_seed(seed)
    protos = torch.rand(10, IMAGE_PIXELS * IMAGE_PIXELS, generator=torch.Generator().manual_seed(7))
    y = torch.randint(0, 10, (n,), generator=g)
    x = (protos[y] + 0.3 * torch.randn(n, IMAGE_PIXELS * IMAGE_PIXELS, generator=g)).clamp(0, 1)
    return x, y


def init_vars(hidden):
    g = torch.Generator().manual_seed(0)
    return {
        "hid_w": torch.randn(IMAGE_PIXELS ** 2, hidden, generator=g) / IMAGE_PIXELS,
        "hid_b": torch.zeros(hidden),
        "sm_w": torch.randn(hidden, 10, generator=g) / hidden ** 0.5,
        "sm_b": torch.zeros(10),
    }


def loss_fn(v, x, y):
    hid = torch.relu(x @ v["hid_w"] + v["hid_b"])
    prob = torch.softmax(hid @ v["sm_w"] + v["sm_b"], dim=1)
    onehot = torch.nn.functional.one_hot(y, 10).float()
    return -(onehot * torch.log(prob.clamp(1e-10, 1.0))).sum()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--train_steps", type=int, default=200)
    ap.add_argument("--batch_size", type=int, default=100)
    ap.add_argument("--hidden_units", type=int, default=100)
    ap.add_argument("--learning_rate", type=float, default=0.01)
    args = ap.parse_args()

    ctx = ReplicaContext.from_env()
    print(f"job name = {ctx.task_type}\ntask index = {ctx.task_index}", flush=True)
    if ctx.kind != "TFJob" or not ctx.cluster.get("ps"):
        # local (non-distributed) run: one in-process PS
        ps = ParameterServer("127.0.0.1", 0, lr=args.learning_rate).start()
        ps_eps = ["127.0.0.1:%d" % ps._srv.server_address[1]]
    else:
        ps_eps = ctx.resolved("ps")
        if ctx.task_type == "ps":
            host, _, port = ctx.my_endpoint().rpartition(":")
            print(f"PS {ctx.task_index} serving on {host}:{port}", flush=True)
            ParameterServer(host, int(port), lr=args.learning_rate).serve_forever()
            return 0

    client = PSClient(ps_eps)
    names = ["hid_w", "hid_b", "sm_w", "sm_b"]
    client.init(init_vars(args.hidden_units))     # first writer wins (the chief's role)
    is_chief = ctx.is_chief or ctx.kind == "local"
    x_all, y_all = synthetic_mnist(6000, seed=100 + ctx.task_index)
    t0 = time.time()
    local_step = 0
    print(f"Worker {ctx.task_index}: Session initialization complete.", flush=True)
    while True:
        i = (local_step * args.batch_size) % (len(x_all) - args.batch_size)
        xb, yb = x_all[i:i + args.batch_size], y_all[i:i + args.batch_size]
        v, step = client.pull(names)
        if step >= args.train_steps:
            break
        for t in v.values():
            t.requires_grad_(True)
        loss = loss_fn(v, xb, yb)
        grads = torch.autograd.grad(loss, [v[n] for n in names])
        step = client.push({n: g for n, g in zip(names, grads)})
        local_step += 1
        if local_step % 50 == 0 or step >= args.train_steps:
            print(f"{time.time():.3f}: Worker {ctx.task_index}: training step {local_step} done "
                  f"(global step: {step}) loss {loss.item() / args.batch_size:.4f}", flush=True)
        if step >= args.train_steps:
            break
    print(f"Training elapsed time: {time.time() - t0:.3f} s", flush=True)
    xv, yv = synthetic_mnist(1000, seed=999)
    v, _ = client.pull(names)
    val = loss_fn(v, xv, yv).item()
    acc = ((torch.relu(xv @ v["hid_w"] + v["hid_b"]) @ v["sm_w"] + v["sm_b"]).argmax(1) == yv).float().mean().item()
    print(f"After {args.train_steps} training step(s), validation cross entropy = {val:g} "
          f"accuracy = {acc:.3f}", flush=True)
    client.close()
    return 0 if (acc > 0.5 or not is_chief) else 3


if __name__ == "__main__":
    sys.exit(main())
