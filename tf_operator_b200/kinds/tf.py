"""TFJob reconciler: TF_CONFIG generation, its own ReconcilePods (adds the Restarting
condition), and the Chief/Master-or-worker-0 status rules.

Reference: pkg/controller.v1/tensorflow/tfjob_controller.go (:351-508 UpdateJobStatus,
:540-573 SetClusterSpec, :586-617 IsMasterRole / IsWorker0Completed, :644-740 ReconcilePods),
tensorflow.go (:51-173 TF_CONFIG), util.go:29-42, pod.go:298-319 isDistributed.
"""
from __future__ import annotations

import threading

import json
import os
from typing import Any, Dict, List, Optional

from ..api import common as c
from ..api import tensorflow as tfapi
from .base import KindReconciler

TF_CONFIG = "TF_CONFIG"
_ALL_TYPES = (tfapi.CHIEF, tfapi.EVALUATOR, tfapi.MASTER, tfapi.PS, tfapi.WORKER)


def contain_chief_or_master_spec(replicas) -> bool:
    return tfapi.CHIEF in replicas or tfapi.MASTER in replicas


def get_port_from_tfjob(job, rtype: str) -> int:
    """Port named tfjob-port on the tensorflow container, else 2222 (util.go:29-42)."""
    port = c.get_port(job, tfapi.INFO, rtype)
    return port if port is not None else tfapi.INFO.port


def is_distributed(job) -> bool:
    replicas = c.replica_specs(job, tfapi.INFO)
    count = 0
    for typ in _ALL_TYPES:
        spec = replicas.get(typ)
        if spec is not None:
            count += 1 if spec.get("replicas") is None else int(spec["replicas"])
    return count != 1


def gen_cluster_spec(job) -> Dict[str, List[str]]:
    cluster: Dict[str, List[str]] = {}
    name, ns = c.name_of(job), c.namespace_of(job)
    domain = os.environ.get(c.ENV_CUSTOM_CLUSTER_DOMAIN, "")
    for rtype, spec in c.replica_specs(job, tfapi.INFO).items():
        rt = rtype.lower()
        port = get_port_from_tfjob(job, rtype)
        hosts = []
        for i in range(int(spec.get("replicas", 1))):
            svc = f"{c.gen_general_name(name, rt, i)}.{ns}.svc"
            if domain:
                svc += "." + domain
            hosts.append(f"{svc}:{port}")
        cluster[rt] = hosts
    return cluster


def _go_json(obj) -> str:
    """encoding/json layout: no spaces, map keys sorted, struct fields in declaration order
    (we build dicts in that order and sort only the maps)."""
    return json.dumps(obj, separators=(",", ":"), ensure_ascii=False)


def gen_tf_config_json(job, rtype: str, index: str) -> str:
    """TF_CONFIG value, byte-for-byte what Go's json.Marshal emits
    (pinned by pkg/controller.v1/tensorflow/pod_test.go:258-280)."""
    i = int(index, 0)
    rt = rtype.lower()
    cluster = gen_cluster_spec(job)
    if (job.get("spec") or {}).get("enableDynamicWorker"):
        worker: Dict[str, str] = {}
        ps: Optional[List[str]] = []
        if rt == tfapi.PS.lower():
            ps = [cluster[rt][i]]
        elif rt == tfapi.WORKER.lower():
            ps = cluster.get(tfapi.PS.lower())  # nil slice -> null when there is no PS
            worker[str(i)] = cluster[rt][i]
        sparse = {"sparseCluster": {"worker": worker, "ps": ps},
                  "task": {"type": rt, "index": i}}
        return _go_json(sparse)
    cfg = {"cluster": {k: cluster[k] for k in sorted(cluster)},
           "task": {"type": rt, "index": i}, "environment": "cloud"}
    return _go_json(cfg)


class TFJobReconciler(KindReconciler):
    module = tfapi
    created_reason = tfapi.REASON_CREATED
    invalid_spec_reason = "InvalidTFJobSpec"

    def set_cluster_spec(self, job, pod_template, rtype: str, index: str) -> None:
        if not is_distributed(job):
            return  # local training jobs get no TF_CONFIG
        value = gen_tf_config_json(job, rtype, index)
        if not value:
            return
        for ctr in (pod_template.get("spec") or {}).get("containers") or []:
            if ctr.get("name") == tfapi.INFO.container:
                env = list(ctr.get("env") or [])
                env.append({"name": TF_CONFIG, "value": value})
                ctr["env"] = env
                break

    def is_master_role(self, replicas, rtype: str, index: int) -> bool:
        if contain_chief_or_master_spec(replicas):
            return tfapi.is_chief_or_master(rtype)
        return rtype == tfapi.WORKER and index == 0

    def on_pod_restart(self, job, status, rtype: str) -> None:
        msg = f"TFJob {c.name_of(job)} is restarting because {rtype} replica(s) failed."
        self.recorder.event(job, "Warning", tfapi.REASON_RESTARTING, msg)
        c.update_job_conditions(status, c.JOB_RESTARTING, tfapi.REASON_RESTARTING, msg)
        if self.metrics is not None:
            self.metrics.jobs_restarted.labels(c.namespace_of(job)).inc()

    def is_worker0_completed(self, job, replicas) -> bool:
        if tfapi.WORKER not in replicas:
            return True
        pods = self.filter_for_replica_type(self._cached_pods(job), tfapi.WORKER.lower())
        slices = self.get_slices(pods, int(replicas[tfapi.WORKER].get("replicas", 1)))
        for index, sl in enumerate(slices):
            if len(sl) == 1 and index == 0:
                pod = sl[0]
                if get_container_exit_code(pod, tfapi.INFO.container) == 0 and \
                        (pod.get("status") or {}).get("phase") == c.POD_SUCCEEDED:
                    return True
        return False

    # The pod list of the reconcile in flight, so update_job_status() does not list again.
    # Workers of one kind share this reconciler object (Manager.start spawns `threadiness`
    # threads per kind), so the list is per THREAD and tagged with the job it belongs to: a
    # thread only ever reads what its own reconcile of that same job stored.
    _tls = threading.local()

    def _cached_pods(self, job):
        uid = (job.get("metadata") or {}).get("uid") or c.key_of(job)
        if getattr(self._tls, "uid", None) == uid and getattr(self._tls, "pods", None) is not None:
            return self._tls.pods
        return self.get_pods_for_job(job)

    def reconcile_jobs(self, job, replicas, status, run_policy) -> None:
        self._tls.uid = (job.get("metadata") or {}).get("uid") or c.key_of(job)
        self._tls.pods = None
        try:
            super().reconcile_jobs(job, replicas, status, run_policy)
        finally:
            self._tls.uid = self._tls.pods = None

    def reconcile_pods(self, job, status, pods, rtype, spec, replicas) -> None:
        self._tls.pods = pods
        super().reconcile_pods(job, status, pods, rtype, spec, replicas)

    def update_job_status(self, job, replicas, status) -> None:
        name, ns = c.name_of(job), c.namespace_of(job)
        worker0_completed = self.is_worker0_completed(job, replicas)
        if status.get("startTime") is None:
            status["startTime"] = c.now_rfc3339()
            ads = c.run_policy(job).get("activeDeadlineSeconds")
            if ads is not None:
                self.workqueue.add_after(c.key_of(job), float(ads))
        success_policy = (job.get("spec") or {}).get("successPolicy") or tfapi.SUCCESS_POLICY_DEFAULT
        for rtype in _ALL_TYPES:
            spec = replicas.get(rtype)
            if spec is None:
                continue
            rs = (status.get("replicaStatuses") or {}).get(rtype) or {}
            succeeded = int(rs.get("succeeded", 0))
            expected = int(spec.get("replicas", 1)) - succeeded
            running = int(rs.get("active", 0))
            failed = int(rs.get("failed", 0))
            if contain_chief_or_master_spec(replicas):
                if tfapi.is_chief_or_master(rtype):
                    if running > 0:
                        c.update_job_conditions(status, c.JOB_RUNNING, tfapi.REASON_RUNNING,
                                                f"TFJob {ns}/{name} is running.")
                    if expected == 0:
                        self._succeed(job, status, f"TFJob {ns}/{name} successfully completed.")
            elif rtype == tfapi.WORKER:
                if expected == 0 or (worker0_completed and
                                     success_policy != tfapi.SUCCESS_POLICY_ALL_WORKERS):
                    self._succeed(job, status, f"TFJob {ns}/{name} successfully completed.")
                elif running > 0:
                    c.update_job_conditions(status, c.JOB_RUNNING, tfapi.REASON_RUNNING,
                                            f"TFJob {ns}/{name} is running.")
            if failed > 0:
                restarting = any(cd.get("type") == c.JOB_RESTARTING
                                 for cd in status.get("conditions") or [])
                if restarting:
                    if self.metrics is not None:   # counter inflation kept (Appendix B.11)
                        self.metrics.jobs_failed.labels(ns).inc()
                else:
                    msg = f"TFJob {ns}/{name} has failed because {failed} {rtype} replica(s) failed."
                    self.recorder.event(job, "Normal", tfapi.REASON_FAILED, msg)
                    if status.get("completionTime") is None:
                        status["completionTime"] = c.now_rfc3339()
                    c.update_job_conditions(status, c.JOB_FAILED, tfapi.REASON_FAILED, msg)
                    if self.metrics is not None:
                        self.metrics.jobs_failed.labels(ns).inc()
        job["status"] = status

    def _succeed(self, job, status, msg: str) -> None:
        self.recorder.event(job, "Normal", tfapi.REASON_SUCCEEDED, msg)
        if status.get("completionTime") is None:
            status["completionTime"] = c.now_rfc3339()
        c.update_job_conditions(status, c.JOB_SUCCEEDED, tfapi.REASON_SUCCEEDED, msg)
        if self.metrics is not None:
            self.metrics.jobs_successful.labels(c.namespace_of(job)).inc()

    def on_job_failed(self, job) -> None:
        if self.metrics is not None:
            self.metrics.jobs_failed.labels(c.namespace_of(job)).inc()


def get_container_exit_code(pod, container: str) -> int:
    code = 0xBEEF
    for cs in (pod.get("status") or {}).get("containerStatuses") or []:
        term = (cs.get("state") or {}).get("terminated")
        if cs.get("name") == container and term is not None:
            code = int(term.get("exitCode", 0))
    return code
