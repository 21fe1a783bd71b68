#!/usr/bin/env python
"""Control-plane scale: N TFJobs (1 PS + 3 Workers each) submitted at once to the API store +
operator, no replica processes - a listener plays the kubelet and marks every created pod Running
at once, so what is timed is the operator alone: watch events -> work queue -> reconcile -> pod and
service creation -> status / condition updates.

The reference states its goal as "O(100) jobs ... running simultaneously in a single cluster"
(docs/design/tf_job_design_doc.md:24) and publishes no measurement (the scalability tests of
CHANGELOG.md:507-508 were requested, never reported). Its legacy operator writes through a client
throttled to 5 QPS / burst 10 (cmd/tf-operator.v1/app/options/options.go:81-82): N jobs x 8 creates
cannot leave it faster than (8 N - 10) / 5 seconds. That bound is printed next to the measurement
and also reproduced by running this operator with the same throttle (``--with-throttle``; slow).

    python benchmarks/control_plane_scale.py --jobs 100 --threadiness 1 4 --out profiles/r2_control_plane_scale.json
"""
import argparse
import json
import os
import statistics
import sys
import threading
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from tf_operator_b200.api import common as c  # noqa: E402
from tf_operator_b200.operator.manager import Manager, ManagerOptions  # noqa: E402
from tf_operator_b200.store.memory import ApiError, ObjectStore  # noqa: E402


def tfjob(name):
    ctr = {"name": "tensorflow", "image": "none:local", "command": ["true"]}

    def spec(n):
        return {"replicas": n, "restartPolicy": "Never", "template": {"spec": {"containers": [ctr]}}}
    return {"apiVersion": "kubeflow.org/v1", "kind": "TFJob",
            "metadata": {"name": name, "namespace": "default"},
            "spec": {"tfReplicaSpecs": {"PS": spec(1), "Worker": spec(3)}}}


def run(n_jobs, threadiness, qps):
    store = ObjectStore()
    mgr = Manager(store, ManagerOptions(enabled_schemes=["TFJob"], threadiness=threadiness, qps=qps, burst=10))
    running_at, created_at = {}, {}
    lock = threading.Lock()

    def fake_kubelet(resource, etype, obj):
        if resource == "pods" and etype == "ADDED":
            # a pod "starts" the moment it exists (off the notifying thread, like a kubelet would)
            def up(pod=obj):
                for _ in range(5):
                    try:
                        p = store.get("pods", c.namespace_of(pod), c.name_of(pod))
                        p.setdefault("status", {}).update(
                            {"phase": c.POD_RUNNING, "containerStatuses": [
                                {"name": "tensorflow", "state": {"running": {}}, "restartCount": 0, "ready": True}]})
                        store.update_status("pods", p)
                        return
                    except ApiError:
                        time.sleep(0.001)
            threading.Thread(target=up, daemon=True).start()
        elif resource == "tfjobs" and etype in ("MODIFIED", "ADDED"):
            conds = (obj.get("status") or {}).get("conditions") or []
            if any(x["type"] == "Running" and x["status"] == "True" for x in conds):
                with lock:
                    running_at.setdefault(c.name_of(obj), time.perf_counter())
    store.add_listener(fake_kubelet)
    mgr.start()
    try:
        t0 = time.perf_counter()
        for i in range(n_jobs):
            name = f"scale-{i}"
            created_at[name] = time.perf_counter()
            store.create("tfjobs", tfjob(name))
        t_submitted = time.perf_counter() - t0
        deadline = time.time() + max(120.0, 3.0 * (8 * n_jobs) / qps if qps else 120.0)
        while time.time() < deadline:
            with lock:
                if len(running_at) == n_jobs:
                    break
            time.sleep(0.002)
        with lock:
            done = dict(running_at)
        lat = sorted((done[k] - created_at[k]) * 1e3 for k in done)
        pods = len(store.list("pods", "default"))
        svcs = len(store.list("services", "default"))
        return {"jobs": n_jobs, "threadiness": threadiness, "throttle_qps": qps or None,
                "jobs_running": len(done), "pods_created": pods, "services_created": svcs,
                "submit_all_s": round(t_submitted, 3),
                "all_running_s": round(max(done.values()) - t0, 3) if len(done) == n_jobs else None,
                "create_to_running_ms": {"p50": round(statistics.median(lat), 2),
                                         "p90": round(lat[int(0.9 * (len(lat) - 1))], 2),
                                         "max": round(lat[-1], 2)} if lat else None,
                "creates_per_s": round((pods + svcs) / (max(done.values()) - t0), 1) if len(done) == n_jobs else None}
    finally:
        mgr.stop()
        store.remove_listener(fake_kubelet)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--jobs", type=int, nargs="+", default=[100])
    ap.add_argument("--threadiness", type=int, nargs="+", default=[1, 4])
    ap.add_argument("--with-throttle", action="store_true",
                    help="also run with the reference's 5 QPS / burst 10 write throttle (takes minutes)")
    ap.add_argument("--out", default="")
    a = ap.parse_args()
    res = {"what": "N TFJobs (1 PS + 3 Workers) created at once; pods marked Running by a listener; "
                   "time until every job shows the Running condition (operator only, CPU)",
           "reference_lower_bound": "5 QPS / burst 10 client throttle: (8 N - 10) / 5 s for N jobs "
                                    "(options.go:81-82); no published measurement",
           "runs": []}
    for n in a.jobs:
        for th in a.threadiness:
            r = run(n, th, 0.0)
            r["reference_throttle_bound_s"] = round((8 * n - 10) / 5.0, 1)
            res["runs"].append(r)
            print(json.dumps(r), flush=True)
        if a.with_throttle:
            r = run(n, 1, 5.0)
            r["reference_throttle_bound_s"] = round((8 * n - 10) / 5.0, 1)
            res["runs"].append(r)
            print(json.dumps(r), flush=True)
    if a.out:
        with open(a.out, "w") as f:
            json.dump(res, f, indent=1)


if __name__ == "__main__":
    main()
