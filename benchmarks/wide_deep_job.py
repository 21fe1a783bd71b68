#!/usr/bin/env python
"""BASELINE.json config 4 through the product, next to its stock comparator.

Submits examples/wide_deep_ps/tfjob.yaml (TFJob, PS + Worker replicas, one GPU each) to a
LocalCluster with ``TFJobClient``, waits for it, and sums the workers' device-timed throughput;
then runs benchmarks/wide_deep_baseline.py (stock PyTorch: NCCL all_to_all pull / push,
torch.unique + index_add_ Adagrad, DDP tower) under torchrun with the same PS / worker split,
table, batch and id distribution. One JSON document with both arms.

    python benchmarks/wide_deep_job.py --ps 2 --workers 6 --out profiles/r2_wide_deep_8gpu.json
"""
import argparse
import json
import os
import subprocess
import sys
import time

import yaml

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)


def ours(args):
    from tf_operator_b200.cluster import LocalCluster
    from tf_operator_b200.launcher.kubelet import KubeletOptions
    from tf_operator_b200.sdk import TFJobClient
    job = yaml.safe_load(open(os.path.join(REPO, "examples", "wide_deep_ps", "tfjob.yaml")))
    extra = [f"--rows={args.rows}", f"--batch-size={args.batch_size}", f"--steps={args.steps}",
             f"--warmup={args.warmup}"]
    specs = job["spec"]["tfReplicaSpecs"]
    specs["PS"]["replicas"], specs["Worker"]["replicas"] = args.ps, args.workers
    for rt in specs:
        ctr = specs[rt]["template"]["spec"]["containers"][0]
        ctr["command"] = list(ctr["command"]) + extra
    cl = LocalCluster(kubelet_options=KubeletOptions(cwd=REPO))
    try:
        cli = TFJobClient(cluster=cl)
        t0 = time.time()
        cli.create(job)
        name = job["metadata"]["name"]
        cli.wait_for_job(name, polling_interval=0.2, timeout_seconds=900)
        ok = cli.is_job_succeeded(name)
        logs = cli.get_logs(name, master=False)
        rows = []
        for pod, text in sorted(logs.items()):
            for ln in text.splitlines():
                if ln.startswith('{"role": "worker"'):
                    rows.append(json.loads(ln))
        if not ok or len(rows) != args.workers:
            raise RuntimeError("job failed: " + "\n".join(f"--- {k}\n{v[-1500:]}" for k, v in logs.items()))
        slowest = max(r["ms_per_step"] for r in rows)
        return {"impl": "tf_operator_b200: TFJob via TFJobClient -> LocalCluster; K4 gather / K5b "
                        "push over NVLink peer memory, tcgen05 tower, fused P2P all-reduce",
                "ps": args.ps, "workers": args.workers, "rows": args.rows,
                "batch_per_worker": args.batch_size,
                "samples_per_s": round(sum(r["samples_per_s_this_worker"] for r in rows), 1),
                "samples_per_s_at_slowest_worker": round(args.workers * args.batch_size / slowest * 1e3, 1),
                "ms_per_step_per_worker": [r["ms_per_step"] for r in rows],
                "kernel_launches_per_step": rows[0]["kernel_launches_per_step"],
                "loss": rows[0]["loss"], "job_wall_s": round(time.time() - t0, 1),
                "note": "workers are asynchronous towards the PS (as in the reference's PS training); "
                        "the dense tower is all-reduced among the workers every step"}
    finally:
        cl.stop()


def stock(args):
    out = os.path.join("/tmp", f"wide_deep_stock_{os.getpid()}.json")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1",
           f"--nproc-per-node={args.ps + args.workers}", "--master-addr", "127.0.0.1",
           "--master-port", str(args.port), os.path.join(REPO, "benchmarks", "wide_deep_baseline.py"),
           f"--ps={args.ps}", f"--rows={args.rows}", f"--batch-size={args.batch_size}",
           f"--steps={args.steps}", f"--warmup={args.warmup}", f"--out={out}"]
    p = subprocess.run(cmd, cwd=REPO, capture_output=True, text=True, timeout=900)
    if p.returncode != 0 or not os.path.exists(out):
        return {"error": (p.stderr or p.stdout)[-1500:]}
    return json.load(open(out))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ps", type=int, default=2)
    ap.add_argument("--workers", type=int, default=6)
    ap.add_argument("--rows", type=int, default=10_000_000)
    ap.add_argument("--batch-size", type=int, default=8192)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--port", type=int, default=29541)
    ap.add_argument("--out", default="")
    ap.add_argument("--skip-stock", action="store_true")
    args = ap.parse_args()
    res = {"config": "BASELINE.json config 4: wide-and-deep, sharded embedding, PS + workers",
           "native": ours(args)}
    if not args.skip_stock:
        res["stock"] = stock(args)
        if "samples_per_s" in res["stock"]:
            res["native_over_stock"] = round(res["native"]["samples_per_s"] / res["stock"]["samples_per_s"], 2)
    line = json.dumps(res, indent=1)
    print(line)
    if args.out:
        with open(args.out, "w") as f:
            f.write(line + "\n")


if __name__ == "__main__":
    main()
