"""Launch latency on REAL GPU replicas, through the product's own API.

``measure(cluster, replicas, trials)`` submits a PyTorchJob of ``replicas`` one-GPU replicas
(examples/launch_probe/probe.py) with ``PyTorchJobClient.create`` on a warm LocalCluster and
reports, per trial and as p50:

* ``launch_to_running_ms``    - ``create`` call -> ``Running`` condition visible through the SDK;
* ``launch_to_first_step_ms`` - ``create`` call -> the instant the replicas' first all-reduce over
  mapped peer memory had completed (spawn + CUDA context + rendezvous + IPC mapping + kernel),
  read from the master replica's log.

BASELINE.json's metric names "job launch-to-Running p50"; the reference can only observe it at its
30 s poll (py/kubeflow/tf_operator/tf_job_client.py:121-122) and creates pods under a 5 QPS /
burst 10 client throttle (cmd/tf-operator.v1/app/options/options.go:81-82).
"""
from __future__ import annotations

import json
import os
import statistics
import time
from typing import Any, Dict

REPO = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def probe_job(name: str, replicas: int) -> Dict[str, Any]:
    ctr = {"name": "pytorch", "image": "tf-operator-b200/launch-probe:local",
           "command": ["python", os.path.join(REPO, "examples", "launch_probe", "probe.py")],
           "resources": {"limits": {"nvidia.com/gpu": 1}}}

    def spec(n):
        return {"replicas": n, "restartPolicy": "Never", "template": {"spec": {"containers": [ctr]}}}
    specs = {"Master": spec(1)}
    if replicas > 1:
        specs["Worker"] = spec(replicas - 1)
    return {"apiVersion": "kubeflow.org/v1", "kind": "PyTorchJob",
            "metadata": {"name": name, "namespace": "default"},
            "spec": {"runPolicy": {"cleanPodPolicy": "None"}, "pytorchReplicaSpecs": specs}}


def measure(cluster, replicas: int, trials: int = 10, timeout: float = 120.0) -> Dict[str, Any]:
    from ..sdk import PyTorchJobClient
    cli = PyTorchJobClient(cluster=cluster)
    running, first, detail = [], [], []
    for t in range(trials):
        name = f"launch-probe-{replicas}-{t}"
        t0 = time.time()
        cli.create(probe_job(name, replicas))
        t_running = None
        while True:
            conds = (cli.get(name).get("status") or {}).get("conditions") or []
            types = {c["type"] for c in conds if c.get("status") == "True"}
            if t_running is None and ("Running" in types or "Succeeded" in types):
                t_running = time.time()
            if conds and conds[-1]["type"] in ("Succeeded", "Failed"):
                break
            if time.time() - t0 > timeout:
                raise TimeoutError(f"{name} did not finish within {timeout} s")
            time.sleep(0.001)
        if not cli.is_job_succeeded(name):
            raise RuntimeError(f"{name} failed: {cli.get_logs(name, master=False)}")
        log = cli.get_logs(name, master=True)[f"{name}-master-0"]
        ev = next(json.loads(ln) for ln in log.splitlines() if ln.startswith('{"event": "first_step"'))
        running.append((t_running - t0) * 1e3)
        first.append((ev["unix_time"] - t0) * 1e3)
        detail.append({"spawn_to_process_start_ms": round((ev["process_start_unix_time"] - t0) * 1e3, 2),
                       "cuda_context_ms": round(ev["cuda_context_s"] * 1e3, 2),
                       "rendezvous_and_map_ms": round(ev["rendezvous_and_map_s"] * 1e3, 2),
                       "first_allreduce_ms": round(ev["first_allreduce_s"] * 1e3, 2)})
        cli.delete(name)
        end = time.time() + 20
        while time.time() < end and any(name in p["metadata"]["name"] for p in cluster.store.list("pods")):
            time.sleep(0.005)
    return {"replicas": replicas, "trials": trials,
            "launch_to_running_ms": round(statistics.median(running), 2),
            "launch_to_first_step_ms": round(statistics.median(first), 2),
            "launch_to_running_ms_all": [round(x, 2) for x in running],
            "launch_to_first_step_ms_all": [round(x, 2) for x in first],
            "breakdown_p50_ms": {k: round(statistics.median(d[k] for d in detail), 2)
                                 for k in detail[0]},
            "what": "PyTorchJobClient.create on a warm LocalCluster -> Running condition / first "
                    "all-reduce over CUDA-IPC-mapped peer memory completed on real one-GPU replicas"}
