#!/usr/bin/env python
"""Single-worker MNIST MLP with training summaries — the reference's mnist_with_summaries
(examples/v1/mnist_with_summaries/mnist_with_summaries.py:36-176): 784->500->10 MLP with
dropout, Adam, scalar summaries (accuracy / cross-entropy / weight statistics) every 10 steps and
a full step trace (wall time per phase) every 100 steps, written as JSON lines under --log_dir."""
import argparse
import json
import os
import sys
import time

import torch


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--max_steps", type=int, default=300)
    ap.add_argument("--learning_rate", type=float, default=0.001)
    ap.add_argument("--dropout", type=float, default=0.9, help="keep probability")
    ap.add_argument("--log_dir", default="/tmp/tfop-train/mnist_with_summaries")
    ap.add_argument("--batch_size", type=int, default=100)
    args = ap.parse_args()
    os.makedirs(os.path.join(args.log_dir, "train"), exist_ok=True)
    os.makedirs(os.path.join(args.log_dir, "test"), exist_ok=True)
    train_w = open(os.path.join(args.log_dir, "train", "events.jsonl"), "a")
    test_w = open(os.path.join(args.log_dir, "test", "events.jsonl"), "a")
    torch.manual_seed(0)
    model = torch.nn.Sequential(torch.nn.Linear(784, 500), torch.nn.ReLU(), torch.nn.Dropout(1 - args.dropout),
                                torch.nn.Linear(500, 10))
    opt = torch.optim.Adam(model.parameters(), lr=args.learning_rate)
    protos = torch.rand(10, 784, generator=torch.Generator().manual_seed(7))
    g = torch.Generator().manual_seed(1)

    def data(n):
        y = torch.randint(0, 10, (n,), generator=g)
        return (protos[y] + 0.3 * torch.randn(n, 784, generator=g)).clamp(0, 1), y

    def summaries(step, x, y, writer, extra=None):
        with torch.no_grad():
            logits = model(x)
            rec = {"step": step, "cross_entropy": torch.nn.functional.cross_entropy(logits, y).item(),
                   "accuracy": (logits.argmax(1) == y).float().mean().item()}
            for i, layer in enumerate(m for m in model if isinstance(m, torch.nn.Linear)):
                w = layer.weight
                rec[f"layer{i + 1}/weights"] = {"mean": w.mean().item(), "stddev": w.std().item(),
                                                "max": w.max().item(), "min": w.min().item()}
        rec.update(extra or {})
        writer.write(json.dumps(rec) + "\n")
        writer.flush()
        return rec

    acc = 0.0
    for i in range(args.max_steps):
        if i % 10 == 0:
            model.eval()
            acc = summaries(i, *data(1000), test_w)["accuracy"]
            print(f"Accuracy at step {i}: {acc}", flush=True)
            model.train()
        x, y = data(args.batch_size)
        t0 = time.perf_counter()
        opt.zero_grad()
        loss = torch.nn.functional.cross_entropy(model(x), y)
        t1 = time.perf_counter()
        loss.backward()
        t2 = time.perf_counter()
        opt.step()
        t3 = time.perf_counter()
        if i % 100 == 99:      # the FULL_TRACE run metadata of the reference
            summaries(i, x, y, train_w, {"trace_ms": {"forward": (t1 - t0) * 1e3, "backward": (t2 - t1) * 1e3,
                                                      "optimizer": (t3 - t2) * 1e3}})
            print(f"Adding run metadata for {i}", flush=True)
    return 0 if acc > 0.8 else 3


if __name__ == "__main__":
    sys.exit(main())
