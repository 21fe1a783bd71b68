"""Behavioural tables for the PyTorchJob, MXJob and XGBoostJob reconcilers. The reference has
NO unit tests for these ("the new reconcilers have no behavioural unit
tests"), so the expectations are read off the reconcilers' source (pytorchjob_controller.go
:315-393, mxjob_controller.go:328-410, xgboostjob_controller.go:324-402) and pinned here with the
same fake-control / hand-filled-store technique as the TFJob tables."""
import pytest

from fixtures import make_controller, new_pod, replica_template, set_pods_statuses, set_services
from tf_operator_b200.api import common as c
from tf_operator_b200.api import mxnet as mxapi
from tf_operator_b200.api import pytorch as ptapi
from tf_operator_b200.api import xgboost as xgbapi
from tf_operator_b200.kinds.mxnet import MXJobReconciler
from tf_operator_b200.kinds.pytorch import PyTorchJobReconciler
from tf_operator_b200.kinds.xgboost import XGBoostJobReconciler


def job_of(kind, specs_key, specs, name="j", **run_policy):
    container, port_name, port = {
        "PyTorchJob": ("pytorch", "pytorchjob-port", 23456),
        "MXJob": ("mxnet", "mxjob-port", 9091),
        "XGBoostJob": ("xgboost", "xgboostjob-port", 9999)}[kind]
    rs = {rt: {"replicas": n, **({"restartPolicy": pol} if pol else {}),
               "template": replica_template(container, port_name, port)}
          for rt, (n, pol) in specs.items()}
    job = {"apiVersion": "kubeflow.org/v1", "kind": kind,
           "metadata": {"name": name, "namespace": "default", "uid": "uid-" + name},
           "spec": {specs_key: rs}}
    if kind == "MXJob":
        job["spec"]["jobMode"] = "MXTrain"
    if run_policy:
        job["spec"]["runPolicy"] = run_policy
    return job


def reconcile(ctrl, job):
    ctrl.module.set_defaults(job)
    status = c.status_of(job)
    ctrl.reconcile_jobs(job, c.replica_specs(job, ctrl.info), status, ctrl.run_policy_for(job)
                        if hasattr(ctrl, "run_policy_for") else c.run_policy(job))
    return status


def last(status):
    conds = status.get("conditions") or []
    return conds[-1]["type"] if conds else None


# ---------------------------------------------------------------------------------- PyTorchJob
PT = [
    # name, master (pending, running, succeeded, failed), worker (...), policy, creates, last condition
    ("fresh job creates master + workers and one service", (0, 0, 0, 0), (0, 0, 0, 0), None, (4, 1), None),
    ("master running -> Running", (0, 1, 0, 0), (0, 3, 0, 0), None, (0, 0), c.JOB_RUNNING),
    ("workers running, master pending -> still Running (trailing update)", (1, 0, 0, 0), (0, 3, 0, 0), None, (0, 0), c.JOB_RUNNING),
    ("master succeeded -> Succeeded even with workers running", (0, 0, 1, 0), (0, 3, 0, 0), None, (0, 0), c.JOB_SUCCEEDED),
    ("worker failed, OnFailure default -> Failed", (0, 1, 0, 0), (0, 2, 0, 1), None, (0, 0), c.JOB_FAILED),
    # reference quirk: the Restarting condition is set, then the trailing unconditional Running
    # update replaces it again (pytorchjob_controller.go:358-366 then :383-390)
    ("worker failed under ExitCode -> Restarting, then overwritten by Running", (0, 1, 0, 0), (0, 2, 0, 1), "ExitCode", (0, 0), c.JOB_RUNNING),
    ("master failed -> Failed", (0, 0, 0, 1), (0, 3, 0, 0), None, (0, 0), c.JOB_FAILED),
]


@pytest.mark.parametrize("name,master,worker,policy,creates,cond", PT, ids=[r[0] for r in PT])
def test_pytorchjob_status_table(name, master, worker, policy, creates, cond):
    job = job_of("PyTorchJob", "pytorchReplicaSpecs", {"Master": (1, policy), "Worker": (3, policy)})
    ctrl, store = make_controller(PyTorchJobReconciler, job)
    set_pods_statuses(store, job, "master", *master)
    set_pods_statuses(store, job, "worker", *worker)
    if sum(master):
        set_services(store, job, "master", 1)
    status = reconcile(ctrl, job)
    assert (len(ctrl.pod_control.templates), len(ctrl.service_control.services)) == creates, name
    if cond is None:
        assert last(status) in (None, c.JOB_RUNNING)      # Created is added by the watch predicate
    else:
        assert last(status) == cond, (name, status.get("conditions"))
    if cond in (c.JOB_SUCCEEDED, c.JOB_FAILED):
        assert status.get("completionTime")


def test_pytorchjob_created_pods_env_names_and_default_policy():
    job = job_of("PyTorchJob", "pytorchReplicaSpecs", {"Master": (1, None), "Worker": (2, None)},
                 name="pt")
    ctrl, store = make_controller(PyTorchJobReconciler, job)
    reconcile(ctrl, job)
    pods = {t["metadata"]["name"]: t for t in ctrl.pod_control.templates}
    assert sorted(pods) == ["pt-master-0", "pt-worker-0", "pt-worker-1"]
    assert pods["pt-master-0"]["metadata"]["labels"][c.LABEL_JOB_ROLE] == "master"
    assert c.LABEL_JOB_ROLE not in pods["pt-worker-0"]["metadata"]["labels"]
    assert all(p["spec"]["restartPolicy"] == "OnFailure" for p in pods.values())
    env = {e["name"]: e["value"] for e in pods["pt-worker-1"]["spec"]["containers"][0]["env"]}
    assert env == {"MASTER_PORT": "23456", "MASTER_ADDR": "pt-master-0", "WORLD_SIZE": "3", "RANK": "2",
                   "PYTHONUNBUFFERED": "0"}
    menv = {e["name"]: e["value"] for e in pods["pt-master-0"]["spec"]["containers"][0]["env"]}
    assert menv["MASTER_ADDR"] == "localhost" and menv["RANK"] == "0"
    # only the master gets a service
    assert [s["metadata"]["name"] for s in ctrl.service_control.services] == ["pt-master-0"]


def test_pytorchjob_exit_code_retryable_worker_is_deleted_for_recreation():
    job = job_of("PyTorchJob", "pytorchReplicaSpecs", {"Master": (1, "ExitCode"), "Worker": (2, "ExitCode")})
    ctrl, store = make_controller(PyTorchJobReconciler, job)
    set_pods_statuses(store, job, "master", 0, 1, 0, 0)
    set_services(store, job, "master", 1)
    for i, (phase, code) in enumerate(((c.POD_RUNNING, None), (c.POD_FAILED, 130))):
        pod = new_pod(job, "worker", i, phase, exit_code=code, container="pytorch")
        created = store.create("pods", pod)
        created["status"] = pod["status"]
        store.update_status("pods", created)
    status = reconcile(ctrl, job)
    assert ctrl.pod_control.delete_pod_name == ["j-worker-1"]
    assert last(status) == c.JOB_RUNNING          # Restarting is overwritten by the trailing update
    assert any(c.REASON_JOB_RESTARTING in e for e in ctrl.recorder.events)
    # a permanent exit code is not retried: nothing deleted, job Failed
    job2 = job_of("PyTorchJob", "pytorchReplicaSpecs", {"Master": (1, "ExitCode"), "Worker": (1, "ExitCode")},
                  name="k")
    ctrl2, store2 = make_controller(PyTorchJobReconciler, job2)
    set_pods_statuses(store2, job2, "master", 0, 1, 0, 0)
    pod = new_pod(job2, "worker", 0, c.POD_FAILED, exit_code=1, container="pytorch")
    created = store2.create("pods", pod)
    created["status"] = pod["status"]
    store2.update_status("pods", created)
    reconcile(ctrl2, job2)
    assert ctrl2.pod_control.delete_pod_name == []


def test_pytorchjob_scheduling_policy_is_dropped_like_the_reference():
    job = job_of("PyTorchJob", "pytorchReplicaSpecs", {"Master": (1, None)},
                 schedulingPolicy={"minAvailable": 1}, cleanPodPolicy="All")
    ctrl, _ = make_controller(PyTorchJobReconciler, job)
    rp = ctrl.run_policy_for(job)
    assert "schedulingPolicy" not in rp and rp["cleanPodPolicy"] == "All"


# ---------------------------------------------------------------------------------- MXJob
MX = [
    ("any type running -> Running", {"scheduler": (0, 1, 0, 0), "server": (1, 0, 0, 0), "worker": (1, 0, 0, 0)}, None, c.JOB_RUNNING),
    ("ANY type fully succeeded -> Succeeded (reference quirk)", {"scheduler": (0, 1, 0, 0), "server": (0, 1, 0, 0), "worker": (0, 0, 1, 0)}, None, c.JOB_SUCCEEDED),
    ("a failed replica -> Failed", {"scheduler": (0, 1, 0, 0), "server": (0, 1, 0, 0), "worker": (0, 0, 0, 1)}, None, c.JOB_FAILED),
    ("failed under ExitCode -> Restarting", {"scheduler": (0, 1, 0, 0), "server": (0, 1, 0, 0), "worker": (0, 0, 0, 1)}, "ExitCode", c.JOB_RESTARTING),
]


@pytest.mark.parametrize("name,pods,policy,cond", MX, ids=[r[0] for r in MX])
def test_mxjob_status_table(name, pods, policy, cond):
    job = job_of("MXJob", "mxReplicaSpecs", {"Scheduler": (1, policy), "Server": (1, policy), "Worker": (1, policy)})
    ctrl, store = make_controller(MXJobReconciler, job)
    for rt, counts in pods.items():
        set_pods_statuses(store, job, rt, *counts)
        set_services(store, job, rt, 1)
    status = reconcile(ctrl, job)
    assert last(status) == cond, (name, status.get("conditions"))


def test_mxjob_master_role_is_server_and_created_message_quirk():
    job = job_of("MXJob", "mxReplicaSpecs", {"Scheduler": (1, None), "Server": (2, None), "Worker": (1, None)},
                 name="mx")
    ctrl, _ = make_controller(MXJobReconciler, job)
    reconcile(ctrl, job)
    roles = {t["metadata"]["name"]: t["metadata"]["labels"].get(c.LABEL_JOB_ROLE) for t in ctrl.pod_control.templates}
    assert roles == {"mx-scheduler-0": None, "mx-server-0": "master", "mx-server-1": "master", "mx-worker-0": None}
    assert "xgboostJob" in ctrl.created_message            # sic, mxjob_controller.go:459
    assert all(t["spec"]["restartPolicy"] == "Never" for t in ctrl.pod_control.templates)
    assert len(ctrl.service_control.services) == 4


# ---------------------------------------------------------------------------------- XGBoostJob
XGB = [
    ("master running -> Running", (0, 1, 0, 0), (0, 2, 0, 0), None, c.JOB_RUNNING),
    ("master succeeded -> Succeeded", (0, 0, 1, 0), (0, 2, 0, 0), None, c.JOB_SUCCEEDED),
    ("worker failed -> Failed", (0, 1, 0, 0), (0, 1, 0, 1), None, c.JOB_FAILED),
    # same trailing Running update as the PyTorch reconciler (xgboostjob_controller.go:392-399)
    ("worker failed under ExitCode -> Restarting, then overwritten by Running", (0, 1, 0, 0), (0, 1, 0, 1), "ExitCode", c.JOB_RUNNING),
]


@pytest.mark.parametrize("name,master,worker,policy,cond", XGB, ids=[r[0] for r in XGB])
def test_xgboostjob_status_table(name, master, worker, policy, cond):
    job = job_of("XGBoostJob", "xgbReplicaSpecs", {"Master": (1, policy), "Worker": (2, policy)})
    ctrl, store = make_controller(XGBoostJobReconciler, job)
    set_pods_statuses(store, job, "master", *master)
    set_pods_statuses(store, job, "worker", *worker)
    set_services(store, job, "master", 1)
    set_services(store, job, "worker", 2)
    status = reconcile(ctrl, job)
    assert last(status) == cond, (name, status.get("conditions"))


def test_xgboostjob_env_and_lightgbm_worker_addrs():
    job = job_of("XGBoostJob", "xgbReplicaSpecs", {"Master": (1, None), "Worker": (2, None)}, name="xg")
    ctrl, _ = make_controller(XGBoostJobReconciler, job)
    reconcile(ctrl, job)
    pods = {t["metadata"]["name"]: t for t in ctrl.pod_control.templates}
    env = {e["name"]: e["value"] for e in pods["xg-worker-1"]["spec"]["containers"][0]["env"]}
    assert env["MASTER_ADDR"] == "xg-master-0" and env["MASTER_PORT"] == "9999"
    assert env["WORLD_SIZE"] == "3" and env["RANK"] == "2" and env["PYTHONUNBUFFERED"] == "0"
    assert env["WORKER_PORT"] == "9999" and env["WORKER_ADDRS"] == "xg-worker-0,xg-worker-1"
    menv = {e["name"]: e["value"] for e in pods["xg-master-0"]["spec"]["containers"][0]["env"]}
    assert menv["MASTER_ADDR"] == "xg-master-0" and menv["RANK"] == "0"     # not localhost, unlike PyTorch
    assert len(ctrl.service_control.services) == 3


@pytest.mark.parametrize("cls,kind,key,policy", [
    (PyTorchJobReconciler, "PyTorchJob", "pytorchReplicaSpecs", "None"),
    (MXJobReconciler, "MXJob", "mxReplicaSpecs", "All"),
    (XGBoostJobReconciler, "XGBoostJob", "xgbReplicaSpecs", "All")])
def test_default_clean_pod_policy_per_kind(cls, kind, key, policy):
    specs = {"Scheduler": (1, None), "Server": (1, None), "Worker": (1, None)} if kind == "MXJob" else \
        {"Master": (1, None), "Worker": (1, None)}
    job = job_of(kind, key, specs)
    ctrl, _ = make_controller(cls, job)
    ctrl.module.set_defaults(job)
    assert c.run_policy(job)["cleanPodPolicy"] == policy
