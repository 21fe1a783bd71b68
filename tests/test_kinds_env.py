"""Environment injection + status rules of the PyTorch / XGBoost / MXNet reconcilers.

The reference has no behavioural tests for these ("the new reconcilers have
no behavioural unit tests"); expectations here are derived from the sources cited per test.
"""
import copy
import json

import pytest

from fixtures import make_controller, new_pod
from tf_operator_b200.api import common as c
from tf_operator_b200.api import mxnet as mxapi
from tf_operator_b200.api import pytorch as ptapi
from tf_operator_b200.api import xgboost as xgapi
from tf_operator_b200.kinds import MXJobReconciler, PyTorchJobReconciler, XGBoostJobReconciler

IMG = "img:latest"


def _job(kind, specs_key, container, replicas, extra_spec=None, name="job"):
    specs = {}
    for rt, n in replicas.items():
        specs[rt] = {"replicas": n, "template": {"spec": {"containers": [
            {"name": container, "image": IMG}, {"name": "sidecar", "image": IMG}]}}}
    job = {"apiVersion": "kubeflow.org/v1", "kind": kind,
           "metadata": {"name": name, "namespace": "default", "uid": "u1"},
           "spec": {specs_key: specs, **(extra_spec or {})}}
    return job


def _reconcile(ctrl, job):
    ctrl.module.set_defaults(job)
    st = c.status_of(job)
    ctrl.reconcile_jobs(job, c.replica_specs(job, ctrl.info), st, ctrl.run_policy_for(job))
    return st


def _env(template, idx=0):
    return [(e["name"], e["value"]) for e in template["spec"]["containers"][idx].get("env", [])]


# pkg/controller.v1/pytorch/pytorch.go:13-68 -------------------------------------------------
def test_pytorch_env_every_container_in_order():
    job = _job("PyTorchJob", "pytorchReplicaSpecs", "pytorch", {"Master": 1, "Worker": 2})
    ctrl, _ = make_controller(PyTorchJobReconciler, job)
    _reconcile(ctrl, job)
    by = {t["metadata"]["name"]: t for t in ctrl.pod_control.templates}
    want_master = [("MASTER_PORT", "23456"), ("MASTER_ADDR", "localhost"), ("WORLD_SIZE", "3"),
                   ("RANK", "0"), ("PYTHONUNBUFFERED", "0")]
    assert _env(by["job-master-0"], 0) == want_master and _env(by["job-master-0"], 1) == want_master
    assert _env(by["job-worker-0"]) == [("MASTER_PORT", "23456"), ("MASTER_ADDR", "job-master-0"),
                                         ("WORLD_SIZE", "3"), ("RANK", "1"), ("PYTHONUNBUFFERED", "0")]
    assert dict(_env(by["job-worker-1"]))["RANK"] == "2"
    assert by["job-master-0"]["metadata"]["labels"]["job-role"] == "master"
    assert "job-role" not in by["job-worker-0"]["metadata"]["labels"]
    assert by["job-master-0"]["spec"]["restartPolicy"] == "OnFailure"
    # only the master gets a service
    assert [s["metadata"]["name"] for s in ctrl.service_control.services] == ["job-master-0"]


def test_pytorch_drops_scheduling_policy():
    """pytorchjob_controller.go:153-159."""
    job = _job("PyTorchJob", "pytorchReplicaSpecs", "pytorch", {"Master": 1},
               {"runPolicy": {"schedulingPolicy": {"minAvailable": 5}, "backoffLimit": 3}})
    ctrl, _ = make_controller(PyTorchJobReconciler, job)
    assert ctrl.run_policy_for(job) == {"backoffLimit": 3}


@pytest.mark.parametrize("pods,policy,want_last,want_types", [
    ({("master", 0): c.POD_RUNNING, ("worker", 0): c.POD_RUNNING}, "OnFailure", c.JOB_RUNNING, {c.JOB_RUNNING}),
    ({("master", 0): c.POD_SUCCEEDED, ("worker", 0): c.POD_RUNNING}, "OnFailure", c.JOB_SUCCEEDED, {c.JOB_SUCCEEDED}),
    ({("master", 0): c.POD_RUNNING, ("worker", 0): c.POD_FAILED}, "OnFailure", c.JOB_FAILED, {c.JOB_FAILED}),
    # master succeeded wins over a failed worker because Master is evaluated first (Appendix B.14)
    ({("master", 0): c.POD_SUCCEEDED, ("worker", 0): c.POD_FAILED}, "Never", c.JOB_SUCCEEDED, {c.JOB_SUCCEEDED}),
])
def test_pytorch_status_rules(pods, policy, want_last, want_types):
    job = _job("PyTorchJob", "pytorchReplicaSpecs", "pytorch", {"Master": 1, "Worker": 1})
    for s in job["spec"]["pytorchReplicaSpecs"].values():
        s["restartPolicy"] = policy
    ctrl, store = make_controller(PyTorchJobReconciler, job)
    for (typ, idx), phase in pods.items():
        store.create("pods", new_pod(job, typ, idx, phase, container="pytorch"))
    st = _reconcile(ctrl, job)
    assert st["conditions"][-1]["type"] == want_last
    assert want_types <= {cd["type"] for cd in st["conditions"]}
    if want_last == c.JOB_SUCCEEDED:
        assert st["conditions"][-1]["message"] == "PyTorchJob job is successfully completed."
        assert st.get("completionTime")


def test_pytorch_exit_code_restart_then_running_quirk():
    """Appendix B.12: with ExitCode a failed pod yields Restarting, and the trailing
    unconditional Running update replaces it; a retryable pod is deleted for re-creation."""
    job = _job("PyTorchJob", "pytorchReplicaSpecs", "pytorch", {"Master": 1, "Worker": 1})
    for s in job["spec"]["pytorchReplicaSpecs"].values():
        s["restartPolicy"] = "ExitCode"
    ctrl, store = make_controller(PyTorchJobReconciler, job)
    store.create("pods", new_pod(job, "master", 0, c.POD_RUNNING, container="pytorch"))
    store.create("pods", new_pod(job, "worker", 0, c.POD_FAILED, exit_code=137, container="pytorch"))
    st = _reconcile(ctrl, job)
    assert ctrl.pod_control.delete_pod_name == ["job-worker-0"]
    assert st["conditions"][-1]["type"] == c.JOB_RUNNING and not c.is_failed(st)


# pkg/controller.v1/xgboost/xgboost.go:17-97 -------------------------------------------------
def test_xgboost_env_with_lightgbm_worker_addrs():
    job = _job("XGBoostJob", "xgbReplicaSpecs", "xgboost", {"Master": 1, "Worker": 2})
    ctrl, _ = make_controller(XGBoostJobReconciler, job)
    _reconcile(ctrl, job)
    by = {t["metadata"]["name"]: t for t in ctrl.pod_control.templates}
    assert _env(by["job-master-0"]) == [
        ("MASTER_PORT", "9999"), ("MASTER_ADDR", "job-master-0"), ("WORLD_SIZE", "3"), ("RANK", "0"),
        ("PYTHONUNBUFFERED", "0"), ("WORKER_PORT", "9999"), ("WORKER_ADDRS", "job-worker-0,job-worker-1")]
    assert dict(_env(by["job-worker-1"]))["RANK"] == "2"
    assert len(ctrl.service_control.services) == 3
    single = _job("XGBoostJob", "xgbReplicaSpecs", "xgboost", {"Master": 1}, name="single")
    ctrl2, _ = make_controller(XGBoostJobReconciler, single)
    _reconcile(ctrl2, single)
    names = [n for n, _ in _env(ctrl2.pod_control.templates[0])]
    assert "WORKER_ADDRS" not in names and "WORKER_PORT" not in names


def test_xgboost_keeps_scheduling_policy_and_created_message():
    job = _job("XGBoostJob", "xgbReplicaSpecs", "xgboost", {"Master": 1},
               {"runPolicy": {"schedulingPolicy": {"minAvailable": 1}}})
    ctrl, _ = make_controller(XGBoostJobReconciler, job)
    assert "schedulingPolicy" in ctrl.run_policy_for(job)
    ctrl.store = None
    ctrl.on_owner_create(job)
    cd = job["status"]["conditions"][0]
    assert cd["reason"] == "XGBoostJobCreated" and cd["message"] == "xgboostJob job is created."


# pkg/controller.v1/mxnet/mxnet.go:29-233 -------------------------------------------------
def test_mxnet_env():
    job = _job("MXJob", "mxReplicaSpecs", "mxnet", {"Scheduler": 1, "Server": 2, "Worker": 2},
               {"jobMode": "MXTrain"})
    job["spec"]["mxReplicaSpecs"]["Worker"]["template"]["metadata"] = {
        "annotations": {"tuner-server-key": "k1"}}
    ctrl, _ = make_controller(MXJobReconciler, job)
    _reconcile(ctrl, job)
    by = {t["metadata"]["name"]: t for t in ctrl.pod_control.templates}
    env = _env(by["job-worker-1"])
    assert [n for n, _ in env] == ["MX_CONFIG", "DMLC_PS_ROOT_PORT", "DMLC_PS_ROOT_URI", "DMLC_NUM_SERVER",
                                   "DMLC_NUM_WORKER", "DMLC_ROLE", "DMLC_USE_KUBERNETES", "DMLC_WORKER_ID"]
    d = dict(env)
    assert d["DMLC_PS_ROOT_PORT"] == "9091" and d["DMLC_PS_ROOT_URI"] == "job-scheduler-0"
    assert d["DMLC_NUM_SERVER"] == "2" and d["DMLC_NUM_WORKER"] == "2" and d["DMLC_ROLE"] == "worker"
    assert d["DMLC_USE_KUBERNETES"] == "1" and d["DMLC_WORKER_ID"] == "1"
    assert d["MX_CONFIG"] == (
        '{"cluster":{"scheduler":[{"url":"job-scheduler-0","port":9091}],'
        '"server":[{"url":"job-server-0","port":9091},{"url":"job-server-1","port":9091}],'
        '"worker":[{"url":"job-worker-0","port":9091},{"url":"job-worker-1","port":9091}]},'
        '"labels":{"scheduler":"","server":"","worker":"k1"},"task":{"type":"worker","index":1}}')
    assert "DMLC_WORKER_ID" not in dict(_env(by["job-server-0"]))
    # job-role=master lands on servers (mxjob_controller.go:444-447)
    assert by["job-server-0"]["metadata"]["labels"].get("job-role") == "master"
    assert "job-role" not in by["job-scheduler-0"]["metadata"]["labels"]


def test_mxnet_status_any_type_fully_succeeded_succeeds_job():
    """Appendix B.8."""
    job = _job("MXJob", "mxReplicaSpecs", "mxnet", {"Scheduler": 1, "Server": 1, "Worker": 1},
               {"jobMode": "MXTrain"})
    ctrl, store = make_controller(MXJobReconciler, job)
    store.create("pods", new_pod(job, "scheduler", 0, c.POD_RUNNING, container="mxnet"))
    store.create("pods", new_pod(job, "server", 0, c.POD_RUNNING, container="mxnet"))
    store.create("pods", new_pod(job, "worker", 0, c.POD_SUCCEEDED, container="mxnet"))
    st = _reconcile(ctrl, job)
    assert c.is_succeeded(st) and c.get_condition(st, c.JOB_SUCCEEDED)["reason"] == "MXJobSucceeded"
    ctrl.store = None
    j2 = _job("MXJob", "mxReplicaSpecs", "mxnet", {"Worker": 1}, {"jobMode": "MXTrain"}, name="j2")
    ctrl.on_owner_create(j2)
    assert j2["status"]["conditions"][0]["message"] == "xgboostJob j2 is created."   # sic
