#!/usr/bin/env python
"""Distributed MNIST with a KVStore — the workload shape of the reference's MXJob example
(examples/mxnet/mxjob_dist_v1.yaml:38-42 runs `train_mnist.py --kv-store dist_device_sync`):
roles and addresses come from the DMLC_* / MX_CONFIG environment the operator injects
(pkg/controller.v1/mxnet/mxnet.go:74-118). The scheduler only brokers start-up, servers hold the
parameters (push = aggregate gradients, pull = fetch weights; `dist_sync`: a server applies the
update once every worker has pushed), workers compute gradients."""
import os
import sys
import threading
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from tf_operator_b200.launcher import netmap  # noqa: E402
from tf_operator_b200.runtime.env import ReplicaContext  # noqa: E402
from tf_operator_b200.runtime.ps import ParameterServer, PSClient  # noqa: E402


def main():
    ctx = ReplicaContext.from_env()
    role = os.environ.get("DMLC_ROLE", "worker")
    n_server = int(os.environ.get("DMLC_NUM_SERVER", "1"))
    n_worker = int(os.environ.get("DMLC_NUM_WORKER", "1"))
    print(f"DMLC_ROLE={role} servers={n_server} workers={n_worker} "
          f"root={os.environ.get('DMLC_PS_ROOT_URI')}:{os.environ.get('DMLC_PS_ROOT_PORT')}", flush=True)
    if role == "scheduler":
        threading.Event().wait()                       # ps-lite's scheduler lives as long as the job
    if role == "server":
        host, _, port = ctx.my_endpoint().rpartition(":")
        ParameterServer(host, int(port), lr=0.01).serve_forever()
        return 0
    servers = ctx.resolved("server")
    kv = PSClient(servers)
    torch.manual_seed(0)
    init = {"fc1_w": torch.randn(784, 128) * 0.05, "fc1_b": torch.zeros(128),
            "fc2_w": torch.randn(128, 10) * 0.05, "fc2_b": torch.zeros(10)}
    kv.init(init)
    names = list(init)
    protos = torch.rand(10, 784, generator=torch.Generator().manual_seed(7))
    g = torch.Generator().manual_seed(int(os.environ.get("DMLC_WORKER_ID", "0")))
    for step in range(150):
        y = torch.randint(0, 10, (64,), generator=g)
        x = (protos[y] + 0.3 * torch.randn(64, 784, generator=g)).clamp(0, 1)
        w, _ = kv.pull(names)
        for t in w.values():
            t.requires_grad_(True)
        logits = torch.relu(x @ w["fc1_w"] + w["fc1_b"]) @ w["fc2_w"] + w["fc2_b"]
        loss = torch.nn.functional.cross_entropy(logits, y)
        grads = torch.autograd.grad(loss, [w[n] for n in names])
        kv.push(dict(zip(names, grads)))
    acc = (logits.argmax(1) == y).float().mean().item()
    print(f"worker {os.environ.get('DMLC_WORKER_ID')} final loss {loss.item():.4f} acc {acc:.3f}", flush=True)
    return 0 if acc > 0.7 else 3


if __name__ == "__main__":
    sys.exit(main())
