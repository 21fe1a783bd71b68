"""MXJob reconciler: MX_CONFIG + DMLC_* (ps-lite / BytePS) environment and the
any-replica-type status rules.

Reference: pkg/controller.v1/mxnet/mxnet.go:29-233 (SetPodEnv, genMXConfig, addBytePSEnv),
mxjob_controller.go (:328-410 UpdateJobStatus, :444-447 IsMasterRole = Server,
:459 the copy-pasted "xgboostJob ... is created." message).
"""
from __future__ import annotations

import json

from ..api import common as c
from ..api import mxnet as mxapi
from .base import KindReconciler

MX_CONFIG = "MX_CONFIG"


def gen_mx_config(job, rtype: str, index: str) -> dict:
    i = int(index, 0)
    specs = c.replica_specs(job, mxapi.INFO)
    cluster, labels = {}, {}
    for rt_canon, spec in specs.items():
        rt = rt_canon.lower()
        port = c.get_port(job, mxapi.INFO, rt_canon)
        if port is None:
            raise ValueError("failed to found the port")
        cluster[rt] = [{"url": c.gen_general_name(c.name_of(job), rt, k), "port": int(port)}
                       for k in range(int(spec.get("replicas", 1)))]
        ann = ((spec.get("template") or {}).get("metadata") or {}).get("annotations") or {}
        labels[rt] = ann.get(mxapi.ANNOTATION_TUNER_SERVER_KEY, "")
    return {"cluster": {k: cluster[k] for k in sorted(cluster)},
            "labels": {k: labels[k] for k in sorted(labels)},
            "task": {"type": rtype, "index": i}}


def _addr(cfg, rtype: str, index: int):
    lst = cfg["cluster"].get(rtype.lower(), [])
    return lst[index] if len(lst) > index else {"url": "", "port": 0}


def set_pod_env(job, pod_template, rtype: str, index: str) -> None:
    cfg = gen_mx_config(job, rtype, index)
    value = json.dumps(cfg, separators=(",", ":"))
    root = _addr(cfg, mxapi.SCHEDULER, 0)
    for ctr in (pod_template.get("spec") or {}).get("containers") or []:
        env = list(ctr.get("env") or [])
        env += [
            {"name": MX_CONFIG, "value": value},
            {"name": "DMLC_PS_ROOT_PORT", "value": str(root["port"])},
            {"name": "DMLC_PS_ROOT_URI", "value": root["url"]},
            {"name": "DMLC_NUM_SERVER", "value": str(len(cfg["cluster"].get("server", [])))},
            {"name": "DMLC_NUM_WORKER", "value": str(len(cfg["cluster"].get("worker", [])))},
            {"name": "DMLC_ROLE", "value": cfg["task"]["type"]},
            {"name": "DMLC_USE_KUBERNETES", "value": "1"},
        ]
        if rtype == mxapi.WORKER.lower():   # BytePS needs a per-worker id
            env.append({"name": "DMLC_WORKER_ID", "value": index})
        ctr["env"] = env


class MXJobReconciler(KindReconciler):
    module = mxapi
    created_reason = mxapi.REASON_CREATED
    created_message = "xgboostJob {name} is created."   # sic (mxjob_controller.go:459)

    def set_cluster_spec(self, job, pod_template, rtype: str, index: str) -> None:
        set_pod_env(job, pod_template, rtype, index)

    def is_master_role(self, replicas, rtype: str, index: int) -> bool:
        return rtype == mxapi.SERVER    # sic (mxjob_controller.go:444-447)

    def update_job_status(self, job, replicas, status) -> None:
        name = c.name_of(job)
        for rtype in sorted(replicas):
            spec = replicas[rtype]
            rs = (status.get("replicaStatuses") or {}).get(rtype) or {}
            expected = int(spec.get("replicas", 1)) - int(rs.get("succeeded", 0))
            running, failed = int(rs.get("active", 0)), int(rs.get("failed", 0))
            if status.get("startTime") is None:
                status["startTime"] = c.now_rfc3339()
                ads = c.run_policy(job).get("activeDeadlineSeconds")
                if ads is not None:
                    self.workqueue.add_after(c.key_of(job), float(ads))
            if running > 0:
                c.update_job_conditions(status, c.JOB_RUNNING, mxapi.REASON_RUNNING,
                                        f"MXJob {name} is running.")
            if expected == 0:
                msg = f"MXJob {name} is successfully completed."
                self.recorder.event(job, "Normal", mxapi.REASON_SUCCEEDED, msg)
                if status.get("completionTime") is None:
                    status["completionTime"] = c.now_rfc3339()
                c.update_job_conditions(status, c.JOB_SUCCEEDED, mxapi.REASON_SUCCEEDED, msg)
            if failed > 0:
                if spec.get("restartPolicy") == c.RESTART_EXIT_CODE:
                    msg = f"mxjob {name} is restarting because {failed} {rtype} replica(s) failed."
                    self.recorder.event(job, "Warning", mxapi.REASON_RESTARTING, msg)
                    c.update_job_conditions(status, c.JOB_RESTARTING, mxapi.REASON_RESTARTING, msg)
                else:
                    msg = f"mxjob {name} is failed because {failed} {rtype} replica(s) failed."
                    self.recorder.event(job, "Normal", mxapi.REASON_FAILED, msg)
                    if status.get("completionTime") is None:
                        status["completionTime"] = c.now_rfc3339()
                    c.update_job_conditions(status, c.JOB_FAILED, mxapi.REASON_FAILED, msg)
