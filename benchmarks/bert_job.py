#!/usr/bin/env python
"""BERT-large as the TFJob of BASELINE.json config 3 (Chief + Workers, one GPU each) through
``TFJobClient`` -> LocalCluster, one run per mode: gradient buckets reduced by the NVLS all-reduce,
the linear layers' weight gradients produced and summed by the fused GEMM -> in-switch reduction
kernel (``--fused-wgrad``), and optionally the all-library twin (``--impl=nccl``: torch modules,
DDP + NCCL). Prints / writes one JSON document with the slowest replica's step time per mode.

    python benchmarks/bert_job.py --world 8 --out profiles/r2_bert_job_8gpu.json [--stock]
"""
import argparse
import json
import os
import sys
import time

import yaml

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--world", type=int, default=8)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=4)
    ap.add_argument("--stock", action="store_true", help="also run the all-library twin")
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    from tf_operator_b200.cluster import LocalCluster
    from tf_operator_b200.launcher.kubelet import KubeletOptions
    from tf_operator_b200.sdk import TFJobClient
    modes = [("native_bucket", []), ("native_fused_wgrad", ["--fused-wgrad"])]
    if args.stock:
        modes.append(("stock_ddp_nccl", ["--impl=nccl"]))
    out = {}
    cl = LocalCluster(kubelet_options=KubeletOptions(cwd=REPO))
    try:
        cli = TFJobClient(cluster=cl)
        for mode, extra in modes:
            job = yaml.safe_load(open(os.path.join(REPO, "examples", "bert_large_dp", "tfjob.yaml")))
            name = job["metadata"]["name"] = "bert-" + mode.replace("_", "-")
            specs = job["spec"]["tfReplicaSpecs"]
            for rt in list(specs):
                ctr = specs[rt]["template"]["spec"]["containers"][0]
                ctr["command"] = [c for c in ctr["command"] if not str(c).startswith("--ckpt")] + \
                    [f"--steps={args.steps}", f"--warmup={args.warmup}"] + extra
            if args.world > 1:
                specs["Worker"]["replicas"] = args.world - 1
            else:
                specs.pop("Worker", None)
            t0 = time.time()
            cli.create(job)
            cli.wait_for_job(name, polling_interval=0.2, timeout_seconds=600)
            logs = cli.get_logs(name, master=False)
            rows = [json.loads(ln) for t in logs.values() for ln in t.splitlines() if ln.startswith('{"role"')]
            if not cli.is_job_succeeded(name) or not rows:
                out[mode] = {"error": "\n".join(f"--- {k}\n{v[-1200:]}" for k, v in logs.items())}
            else:
                out[mode] = {"ms_per_step": max(r["ms_per_step"] for r in rows),
                             "seq_per_s_job": min(r["seq_per_s_job"] for r in rows),
                             "fused_wgrad": rows[0].get("fused_wgrad"), "allreduce": rows[0]["allreduce"],
                             "loss": rows[0]["loss"], "kernel_launches_per_step": rows[0]["kernel_launches_per_step"],
                             "world": rows[0]["world"], "job_wall_s": round(time.time() - t0, 1)}
            cli.delete(name)
            time.sleep(2)
    finally:
        cl.stop()
    line = json.dumps(out, indent=1)
    print(line)
    if args.out:
        with open(args.out, "w") as f:
            f.write(line + "\n")


if __name__ == "__main__":
    main()
