"""PyTorchJob reconciler: MASTER_ADDR / MASTER_PORT / WORLD_SIZE / RANK injection and the
Master-driven status rules.

Reference: pkg/controller.v1/pytorch/pytorch.go:13-96 (SetPodEnv),
pytorchjob_controller.go (:153-159 RunPolicy without schedulingPolicy, :315-393
UpdateJobStatus, :434-437 IsMasterRole).
"""
from __future__ import annotations

from ..api import common as c
from ..api import pytorch as ptapi
from .base import KindReconciler


def append_env(pod_template, pairs) -> None:
    """Append (name, value) pairs to EVERY container, in order."""
    for ctr in (pod_template.get("spec") or {}).get("containers") or []:
        env = list(ctr.get("env") or [])
        env.extend({"name": n, "value": v} for n, v in pairs)
        ctr["env"] = env


def set_pod_env(job, pod_template, rtype: str, index: str) -> None:
    rank = int(index)
    specs = c.replica_specs(job, ptapi.INFO)
    total = sum(int(s.get("replicas", 1)) for s in specs.values())
    master_port = c.get_port(job, ptapi.INFO, ptapi.MASTER)
    if master_port is None:
        raise ValueError("port not found")
    master_addr = c.gen_general_name(c.name_of(job), ptapi.MASTER.lower(), 0)
    if rtype == ptapi.MASTER.lower():
        if rank != 0:
            raise ValueError("invalid config: There should be only a single master with index=0")
        master_addr = "localhost"
    else:
        rank += 1
    append_env(pod_template, [
        ("MASTER_PORT", str(master_port)), ("MASTER_ADDR", master_addr),
        ("WORLD_SIZE", str(total)), ("RANK", str(rank)), ("PYTHONUNBUFFERED", "0")])


class PyTorchJobReconciler(KindReconciler):
    module = ptapi
    created_reason = ptapi.REASON_CREATED

    def set_cluster_spec(self, job, pod_template, rtype: str, index: str) -> None:
        set_pod_env(job, pod_template, rtype, index)

    def is_master_role(self, replicas, rtype: str, index: int) -> bool:
        return rtype == ptapi.MASTER

    def services_for(self, rtype: str) -> bool:
        # only the Master needs an address: workers get no default port (defaults.go:98-105)
        return rtype == ptapi.MASTER

    def run_policy_for(self, job):
        rp = dict(c.run_policy(job))
        rp.pop("schedulingPolicy", None)       # pytorchjob_controller.go:153-159
        return rp

    def update_job_status(self, job, replicas, status) -> None:
        name = c.name_of(job)
        if status.get("startTime") is None:
            status["startTime"] = c.now_rfc3339()
            ads = c.run_policy(job).get("activeDeadlineSeconds")
            if ads is not None:
                self.workqueue.add_after(c.key_of(job), float(ads))
        # Go iterates a map here (order undefined); we fix Master first
        for rtype in sorted(replicas, key=lambda t: (t != ptapi.MASTER, t)):
            spec = replicas[rtype]
            rs = (status.get("replicaStatuses") or {}).get(rtype) or {}
            expected = int(spec.get("replicas", 1)) - int(rs.get("succeeded", 0))
            running, failed = int(rs.get("active", 0)), int(rs.get("failed", 0))
            if rtype == ptapi.MASTER:
                if running > 0:
                    c.update_job_conditions(status, c.JOB_RUNNING, c.REASON_JOB_RUNNING,
                                            f"PyTorchJob {name} is running.")
                if expected == 0:
                    msg = f"PyTorchJob {name} is successfully completed."
                    self.recorder.event(job, "Normal", c.REASON_JOB_SUCCEEDED, msg)
                    if status.get("completionTime") is None:
                        status["completionTime"] = c.now_rfc3339()
                    c.update_job_conditions(status, c.JOB_SUCCEEDED, c.REASON_JOB_SUCCEEDED, msg)
                    return
            if failed > 0:
                if spec.get("restartPolicy") == c.RESTART_EXIT_CODE:
                    msg = f"PyTorchJob {name} is restarting because {failed} {rtype} replica(s) failed."
                    self.recorder.event(job, "Warning", c.REASON_JOB_RESTARTING, msg)
                    c.update_job_conditions(status, c.JOB_RESTARTING, c.REASON_JOB_RESTARTING, msg)
                else:
                    msg = f"PyTorchJob {name} is failed because {failed} {rtype} replica(s) failed."
                    self.recorder.event(job, "Normal", c.REASON_JOB_FAILED, msg)
                    if status.get("completionTime") is None:
                        status["completionTime"] = c.now_rfc3339()
                    c.update_job_conditions(status, c.JOB_FAILED, c.REASON_JOB_FAILED, msg)
        # trailing unconditional Running update of the reference (:383-390); a no-op once Failed
        c.update_job_conditions(status, c.JOB_RUNNING, c.REASON_JOB_RUNNING,
                                f"PyTorchJob {name} is running.")
