"""XGBoostJob reconciler (also serves LightGBM): rabit-style rendezvous env.

Reference: pkg/controller.v1/xgboost/xgboost.go:17-135 (SetPodEnv),
xgboostjob_controller.go (:324-402 UpdateJobStatus, :168 full RunPolicy).
"""
from __future__ import annotations

from ..api import common as c
from ..api import xgboost as xgapi
from .base import KindReconciler
from .pytorch import append_env


def set_pod_env(job, pod_template, rtype: str, index: str) -> None:
    rank = int(index)
    specs = c.replica_specs(job, xgapi.INFO)
    if rtype.lower() == xgapi.WORKER.lower():
        rank += int((specs.get(xgapi.MASTER) or {}).get("replicas", 1))
    master_addr = c.gen_general_name(c.name_of(job), xgapi.MASTER.lower(), 0)
    master_port = c.get_port(job, xgapi.INFO, xgapi.MASTER)
    if master_port is None:
        raise ValueError("failed to found the port")
    total = sum(int(s["replicas"]) for s in specs.values() if s and s.get("replicas") is not None)
    pairs = [("MASTER_PORT", str(master_port)), ("MASTER_ADDR", master_addr),
             ("WORLD_SIZE", str(total)), ("RANK", str(rank)), ("PYTHONUNBUFFERED", "0")]
    if total > 1:
        worker_port = c.get_port(job, xgapi.INFO, xgapi.WORKER)
        if worker_port is None:
            raise ValueError("failed to found the port")
        addrs = [c.gen_general_name(c.name_of(job), xgapi.WORKER.lower(), i)
                 for i in range(total - 1)]
        pairs += [("WORKER_PORT", str(worker_port)), ("WORKER_ADDRS", ",".join(addrs))]
    append_env(pod_template, pairs)


class XGBoostJobReconciler(KindReconciler):
    module = xgapi
    created_reason = xgapi.REASON_CREATED
    created_message = "xgboostJob {name} is created."

    def set_cluster_spec(self, job, pod_template, rtype: str, index: str) -> None:
        set_pod_env(job, pod_template, rtype, index)

    def is_master_role(self, replicas, rtype: str, index: int) -> bool:
        return rtype == xgapi.MASTER

    def update_job_status(self, job, replicas, status) -> None:
        name = c.name_of(job)
        if status.get("startTime") is None:
            status["startTime"] = c.now_rfc3339()
            ads = c.run_policy(job).get("activeDeadlineSeconds")
            if ads is not None:
                self.workqueue.add_after(c.key_of(job), float(ads))
        for rtype in sorted(replicas, key=lambda t: (t != xgapi.MASTER, t)):
            spec = replicas[rtype]
            rs = (status.get("replicaStatuses") or {}).get(rtype) or {}
            expected = int(spec.get("replicas", 1)) - int(rs.get("succeeded", 0))
            running, failed = int(rs.get("active", 0)), int(rs.get("failed", 0))
            if rtype == xgapi.MASTER:
                if running > 0:
                    c.update_job_conditions(status, c.JOB_RUNNING, xgapi.REASON_RUNNING,
                                            f"XGBoostJob {name} is running.")
                if expected == 0:
                    msg = f"XGBoostJob {name} is successfully completed."
                    self.recorder.event(job, "Normal", xgapi.REASON_SUCCEEDED, msg)
                    if status.get("completionTime") is None:
                        status["completionTime"] = c.now_rfc3339()
                    c.update_job_conditions(status, c.JOB_SUCCEEDED, xgapi.REASON_SUCCEEDED, msg)
                    return
            if failed > 0:
                if spec.get("restartPolicy") == c.RESTART_EXIT_CODE:
                    msg = f"XGBoostJob {name} is restarting because {failed} {rtype} replica(s) failed."
                    self.recorder.event(job, "Warning", xgapi.REASON_RESTARTING, msg)
                    c.update_job_conditions(status, c.JOB_RESTARTING, xgapi.REASON_RESTARTING, msg)
                else:
                    msg = f"XGBoostJob {name} is failed because {failed} {rtype} replica(s) failed."
                    self.recorder.event(job, "Normal", xgapi.REASON_FAILED, msg)
                    if status.get("completionTime") is None:
                        status["completionTime"] = c.now_rfc3339()
                    c.update_job_conditions(status, c.JOB_FAILED, xgapi.REASON_FAILED, msg)
        c.update_job_conditions(status, c.JOB_RUNNING, xgapi.REASON_RUNNING,
                                f"XGBoostJob {name} is running.")
