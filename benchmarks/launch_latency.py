#!/usr/bin/env python
"""K9: job launch-to-Running latency sweep (BASELINE.json config 5, second half).

For 1/2/4/8-replica gang-scheduled PyTorchJobs: time from the SDK ``create`` call to the
``Running`` condition being visible through the SDK, p50/p90 over trials, for
  * ours      : the operator as shipped (no client throttle), and
  * reference : the same reconcile semantics with the reference operator's client-go write
                throttle (QPS 5, burst 10 - cmd/tf-operator.v1/app/options/options.go:81-82),
                the stand-in for "the reference operator" since it cannot run here
                (BASELINE.md section 2-3).
CPU-side measurement; replicas are `sleep` processes, GPUs are synthetic slots.
"""
import argparse
import json
import os
import statistics
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from tf_operator_b200.cluster import LocalCluster  # noqa: E402
from tf_operator_b200.launcher.kubelet import KubeletOptions  # noqa: E402
from tf_operator_b200.launcher.topology import synthetic  # noqa: E402
from tf_operator_b200.operator.manager import ManagerOptions  # noqa: E402
from tf_operator_b200.sdk import PyTorchJobClient  # noqa: E402


def job(name, replicas):
    ctr = {"name": "pytorch", "image": "sleep:local", "command": ["sleep", "30"],
           "resources": {"limits": {"nvidia.com/gpu": 1}}}
    specs = {"Master": {"replicas": 1, "restartPolicy": "Never", "template": {"spec": {"containers": [ctr]}}}}
    if replicas > 1:
        specs["Worker"] = {"replicas": replicas - 1, "restartPolicy": "Never",
                           "template": {"spec": {"containers": [ctr]}}}
    return {"apiVersion": "kubeflow.org/v1", "kind": "PyTorchJob",
            "metadata": {"name": name, "namespace": "default"},
            "spec": {"runPolicy": {"cleanPodPolicy": "All"}, "pytorchReplicaSpecs": specs}}


def sweep(qps, trials, sizes):
    out = {}
    for n in sizes:
        lat = []
        for t in range(trials):
            # XGBoost/TF honour schedulingPolicy; PyTorchJob drops it but gang pods still form a group
            cl = LocalCluster(topology=synthetic(8), manager_options=ManagerOptions(
                enabled_schemes=["PyTorchJob"], enable_gang_scheduling=True, qps=qps, burst=10))
            try:
                cli = PyTorchJobClient(cluster=cl)
                name = f"lat-{n}-{t}"
                t0 = time.perf_counter()
                cli.create(job(name, n))
                while True:
                    j = cli.get(name)
                    conds = (j.get("status") or {}).get("conditions") or []
                    if any(c["type"] == "Running" and c["status"] == "True" for c in conds):
                        break
                    if time.perf_counter() - t0 > 60:
                        raise TimeoutError(name)
                    time.sleep(0.0005)
                lat.append((time.perf_counter() - t0) * 1e3)
                cli.delete(name)
            finally:
                cl.stop()
        lat.sort()
        out[str(n)] = {"p50_ms": round(statistics.median(lat), 2),
                       "p90_ms": round(lat[min(len(lat) - 1, int(0.9 * len(lat)))], 2),
                       "min_ms": round(lat[0], 2), "trials": trials}
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--trials", type=int, default=10)
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    sizes = [1, 2, 4, 8]
    res = {"metric": "job launch-to-Running latency (SDK create -> Running condition), gang-scheduled",
           "ours": sweep(0.0, args.trials, sizes),
           "reference_semantics_qps5_burst10": sweep(5.0, max(3, args.trials // 3), sizes)}
    print(json.dumps(res, indent=1))
    if args.out:
        json.dump(res, open(args.out, "w"), indent=1)


if __name__ == "__main__":
    main()
