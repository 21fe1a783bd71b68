"""What ``tf.estimator.RunConfig`` derives from TF_CONFIG, without TensorFlow.

The reference's e2e suite checks these fields per replica through test-server's /runconfig
(py/kubeflow/tf_operator/estimator_runconfig_tests.py:25-101): ``master`` is
``grpc://<own endpoint>``, ``num_worker_replicas`` counts chief/master + workers,
``num_ps_replicas`` counts PS, ``is_chief`` follows the chief-or-worker-0 rule, and an
evaluator sees an empty cluster.
"""
from __future__ import annotations

import json
from typing import Any, Dict


def run_config_from_tf_config(tf_config: str) -> Dict[str, Any]:
    if not tf_config:
        return {"task_type": "worker", "task_id": 0, "cluster_spec": {}, "is_chief": True,
                "master": "", "num_worker_replicas": 1, "num_ps_replicas": 0}
    cfg = json.loads(tf_config)
    cluster = cfg.get("cluster") or {}
    task = cfg.get("task") or {}
    ttype, tid = task.get("type", "worker"), int(task.get("index", 0))
    if ttype == "evaluator":
        return {"task_type": "evaluator", "task_id": tid, "cluster_spec": {}, "is_chief": False,
                "master": "", "num_worker_replicas": 0, "num_ps_replicas": 0}
    has_chief = "chief" in cluster or "master" in cluster
    is_chief = ttype in ("chief", "master") if has_chief else (ttype == "worker" and tid == 0)
    members = cluster.get(ttype) or []
    own = members[tid] if tid < len(members) else None
    spec = {k: v for k, v in cluster.items() if k != "evaluator"}
    workers = len(cluster.get("worker") or []) + len(cluster.get("chief") or []) + \
        len(cluster.get("master") or [])
    return {"task_type": ttype, "task_id": tid, "cluster_spec": spec, "is_chief": is_chief,
            "master": f"grpc://{own}" if own else "", "num_worker_replicas": workers,
            "num_ps_replicas": len(cluster.get("ps") or [])}
