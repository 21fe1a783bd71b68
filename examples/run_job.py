#!/usr/bin/env python
"""Submit a job manifest to a fresh single-box cluster, wait for it, print the logs.

    python examples/run_job.py examples/resnet50_ddp/pytorchjob.yaml [--replicas Worker=1] [--timeout 600]

The equivalent of `kubectl apply -f job.yaml` + waiting on the job in the reference's docs
(docs/quick-start-v1.md).
"""
import argparse
import json
import os
import sys
import time

import yaml

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from tf_operator_b200.cluster import LocalCluster  # noqa: E402
from tf_operator_b200.launcher.kubelet import KubeletOptions  # noqa: E402
from tf_operator_b200.operator.manager import ManagerOptions  # noqa: E402
from tf_operator_b200.sdk.client import JobClient  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("manifest")
    ap.add_argument("--replicas", action="append", default=[], help="Type=N overrides")
    ap.add_argument("--arg", action="append", default=[], help="extra container argument")
    ap.add_argument("--timeout", type=float, default=900)
    ap.add_argument("--gang", action="store_true")
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    job = yaml.safe_load(open(args.manifest))
    from tf_operator_b200.api import module_for
    info = module_for(job["kind"]).INFO
    for ov in args.replicas:
        t, n = ov.split("=")
        if int(n) == 0:
            job["spec"][info.specs_key].pop(t, None)
        else:
            job["spec"][info.specs_key][t]["replicas"] = int(n)
    for spec in job["spec"][info.specs_key].values():
        for ctr in spec["template"]["spec"]["containers"]:
            ctr["command"] = list(ctr.get("command", [])) + args.arg
    repo = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    t_create = time.time()
    with LocalCluster(kubelet_options=KubeletOptions(cwd=repo),
                      manager_options=ManagerOptions(enable_gang_scheduling=args.gang)) as cl:

        class _C(JobClient):
            KIND = job["kind"]
        cli = _C(cluster=cl)
        name = job["metadata"]["name"]
        cli.create(job)
        running_at = None
        end = time.time() + args.timeout
        status = ""
        while time.time() < end:
            j = cli.get(name)
            conds = (j.get("status") or {}).get("conditions") or []
            types = [c["type"] for c in conds if c["status"] == "True"]
            if running_at is None and any(c["type"] == "Running" for c in conds):
                running_at = time.time() - t_create
            if "Succeeded" in types or "Failed" in types:
                status = types[-1]
                break
            time.sleep(0.05)
        # read the log files directly: with cleanPodPolicy All/Running the pods may be gone already
        logs = {}
        log_root = cl.kubelet._log_dir / job["metadata"].get("namespace", "default")
        for pod_dir in sorted(log_root.glob(name + "-*")):
            logs[pod_dir.name] = "".join(f.read_text(errors="replace") for f in sorted(pod_dir.glob("*.log")))
        result = {"job": name, "kind": job["kind"], "status": status or "Timeout",
                  "launch_to_running_s": running_at, "wall_s": round(time.time() - t_create, 2)}
        print(json.dumps(result))
        if args.out:
            with open(args.out, "w") as f:
                json.dump({"result": result, "logs": logs}, f, indent=1)
        return 0 if status == "Succeeded" else 1


if __name__ == "__main__":
    sys.exit(main())
