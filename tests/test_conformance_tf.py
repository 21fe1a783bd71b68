"""TFJob conformance: the reference's Go unit-test TABLES re-expressed against our engine
(SURVEY.md section 4 take-away (a)). Each test names the Go test it mirrors."""
import copy
import os
import time

import pytest

from fixtures import (TEST_JOB_NAME, make_controller, new_pod, new_tfjob, set_pods_statuses,
                      set_services)
from tf_operator_b200.api import common as c
from tf_operator_b200.api import tensorflow as tfapi
from tf_operator_b200.engine import expectations as exp
from tf_operator_b200.kinds.tf import (TFJobReconciler, gen_tf_config_json, is_distributed)


def _reconcile(ctrl, job):
    tfapi.set_defaults(job)
    status = c.status_of(job)
    ctrl.reconcile_jobs(job, c.replica_specs(job, ctrl.info), status, c.run_policy(job))
    return status


# controller_test.go:68-198 TestNormalPath ------------------------------------------------
NORMAL_PATH = {
    "Local TFJob is created": (1, 0, (0, 0, 0, 0), (0, 0, 0, 0), (0, 0), (1, 0, 1), (0, 0, 0), (0, 0, 0), None),
    "Distributed TFJob (4 workers, 2 PS) is created": (4, 2, (0, 0, 0, 0), (0, 0, 0, 0), (0, 0), (6, 0, 6), (0, 0, 0), (0, 0, 0), None),
    "all replicas are pending": (4, 2, (4, 0, 0, 0), (2, 0, 0, 0), (4, 2), (0, 0, 0), (0, 0, 0), (0, 0, 0), None),
    "all replicas are running": (4, 2, (0, 4, 0, 0), (0, 2, 0, 0), (4, 2), (0, 0, 0), (4, 0, 0), (2, 0, 0), (c.JOB_RUNNING, tfapi.REASON_RUNNING)),
    "2 workers, 1 PS are pending": (4, 2, (2, 0, 0, 0), (1, 0, 0, 0), (2, 1), (3, 0, 3), (0, 0, 0), (0, 0, 0), None),
    "2 workers, 1 PS are pending, 1 worker is running": (4, 2, (2, 1, 0, 0), (1, 0, 0, 0), (3, 1), (2, 0, 2), (1, 0, 0), (0, 0, 0), (c.JOB_RUNNING, tfapi.REASON_RUNNING)),
    "2 workers, 1 PS are pending, 1 worker is succeeded": (4, 2, (2, 0, 1, 0), (1, 0, 0, 0), (3, 1), (2, 0, 2), (0, 1, 0), (0, 0, 0), None),
    "Distributed TFJob (4 workers, 2 PS) is succeeded": (4, 2, (0, 0, 4, 0), (0, 0, 2, 0), (4, 2), (0, 0, 0), (0, 4, 0), (0, 2, 0), (c.JOB_SUCCEEDED, tfapi.REASON_SUCCEEDED)),
}


@pytest.mark.parametrize("name", list(NORMAL_PATH))
def test_normal_path(name):
    worker, ps, wpods, pspods, svcs, expect, wstat, psstat, cond = NORMAL_PATH[name]
    job = new_tfjob(worker, ps)
    ctrl, store = make_controller(TFJobReconciler, job)
    set_pods_statuses(store, job, "worker", *wpods)
    set_pods_statuses(store, job, "ps", *pspods)
    set_services(store, job, "worker", svcs[0])
    set_services(store, job, "ps", svcs[1])
    status = _reconcile(ctrl, job)
    pc, sc = ctrl.pod_control, ctrl.service_control
    assert len(pc.templates) == expect[0], name
    assert len(pc.delete_pod_name) == expect[1], name
    assert len(sc.services) == expect[2], name
    # every created pod carries the controller ownerRef (controller_test.go:271-287)
    assert len(pc.controller_refs) == expect[0]
    for ref in pc.controller_refs:
        assert ref["apiVersion"] == "kubeflow.org/v1" and ref["kind"] == "TFJob"
        assert ref["name"] == TEST_JOB_NAME and ref["controller"] is True
        assert ref["blockOwnerDeletion"] is True and ref["uid"] == job["metadata"]["uid"]
    rs = status["replicaStatuses"]
    w = rs.get("Worker", {})
    assert (w.get("active", 0), w.get("succeeded", 0), w.get("failed", 0)) == wstat, name
    if ps:
        p = rs.get("PS", {})
        assert (p.get("active", 0), p.get("succeeded", 0), p.get("failed", 0)) == psstat, name
    if cond is not None:
        got = c.get_condition(status, cond[0])
        assert got is not None and got["status"] == "True" and got["reason"] == cond[1], name
    if wpods[1] or pspods[1]:
        assert status.get("startTime")


# job_test.go:191-367 TestDeletePodsAndServices ------------------------------------------------
@pytest.mark.parametrize("policy,wpods,pspods,deletions", [
    (c.CLEAN_ALL, (0, 4, 0, 0), (0, 2, 0, 0), 6),
    (c.CLEAN_RUNNING, (0, 4, 0, 0), (0, 2, 0, 0), 6),
    (c.CLEAN_RUNNING, (0, 0, 4, 0), (0, 0, 2, 0), 0),
    (c.CLEAN_NONE, (0, 0, 4, 0), (0, 0, 2, 0), 0),
    (c.CLEAN_RUNNING, (2, 2, 0, 0), (0, 0, 2, 0), 4),   # Pending counts as Running
    (c.CLEAN_ALL, (0, 0, 4, 0), (0, 0, 2, 0), 6),
])
def test_delete_pods_and_services(policy, wpods, pspods, deletions):
    job = new_tfjob(4, 2, cleanPodPolicy=policy)
    ctrl, store = make_controller(TFJobReconciler, job)
    set_pods_statuses(store, job, "worker", *wpods)
    set_pods_statuses(store, job, "ps", *pspods)
    set_services(store, job, "worker", 4)
    set_services(store, job, "ps", 2)
    tfapi.set_defaults(job)
    status = c.status_of(job)
    c.update_job_conditions(status, c.JOB_SUCCEEDED, tfapi.REASON_SUCCEEDED, "")
    ctrl.reconcile_jobs(job, c.replica_specs(job, ctrl.info), status, c.run_policy(job))
    assert len(ctrl.pod_control.delete_pod_name) == deletions
    assert len(ctrl.service_control.delete_service_name) == deletions
    # the same-named service goes with each pod
    assert sorted(ctrl.pod_control.delete_pod_name) == sorted(ctrl.service_control.delete_service_name)


# job_test.go:369-547 TestCleanupTFJob (commented out upstream; the TTL rule itself) --------------
def test_cleanup_job_ttl():
    job = new_tfjob(1, 0, ttlSecondsAfterFinished=0)
    ctrl, store = make_controller(TFJobReconciler, job)
    tfapi.set_defaults(job)
    status = c.status_of(job)
    status["completionTime"] = c.now_rfc3339()
    c.update_job_conditions(status, c.JOB_SUCCEEDED, tfapi.REASON_SUCCEEDED, "")
    ctrl.reconcile_jobs(job, c.replica_specs(job, ctrl.info), status, c.run_policy(job))
    assert store.list("tfjobs") == []       # TTL 0: deleted immediately
    # a long TTL re-queues instead of deleting
    job2 = new_tfjob(1, 0, ttlSecondsAfterFinished=3600, name="later")
    ctrl2, store2 = make_controller(TFJobReconciler, job2)
    added = []
    ctrl2.workqueue.add_after = lambda k, d: added.append((k, d))
    tfapi.set_defaults(job2)
    st2 = c.status_of(job2)
    st2["completionTime"] = c.now_rfc3339()
    c.update_job_conditions(st2, c.JOB_SUCCEEDED, tfapi.REASON_SUCCEEDED, "")
    ctrl2.reconcile_jobs(job2, c.replica_specs(job2, ctrl2.info), st2, c.run_policy(job2))
    assert len(store2.list("tfjobs")) == 1 and added and 3500 < added[0][1] <= 3600
    # no TTL: nothing happens
    job3 = new_tfjob(1, 0, name="never")
    ctrl3, store3 = make_controller(TFJobReconciler, job3)
    tfapi.set_defaults(job3)
    st3 = c.status_of(job3)
    c.update_job_conditions(st3, c.JOB_SUCCEEDED, tfapi.REASON_SUCCEEDED, "")
    ctrl3.reconcile_jobs(job3, c.replica_specs(job3, ctrl3.info), st3, c.run_policy(job3))
    assert len(store3.list("tfjobs")) == 1


# job_test.go:549-689 TestActiveDeadlineSeconds ------------------------------------------------
def test_active_deadline_seconds():
    job = new_tfjob(4, 2, activeDeadlineSeconds=1)
    ctrl, store = make_controller(TFJobReconciler, job)
    set_pods_statuses(store, job, "worker", 0, 4, 0, 0)
    set_pods_statuses(store, job, "ps", 0, 2, 0, 0)
    set_services(store, job, "worker", 4)
    set_services(store, job, "ps", 2)
    status = _reconcile(ctrl, job)           # first pass sets startTime, job is Running
    assert status.get("startTime") and not c.is_failed(status)
    time.sleep(1.1)
    ctrl.pod_control.clear()
    ctrl.service_control.clear()
    ctrl.reconcile_jobs(job, c.replica_specs(job, ctrl.info), status, c.run_policy(job))
    assert len(ctrl.pod_control.delete_pod_name) == 6
    assert len(ctrl.service_control.delete_service_name) == 6
    failed = c.get_condition(status, c.JOB_FAILED)
    assert failed and failed["reason"] == "JobFailed"
    assert "active longer than specified deadline" in failed["message"]
    assert status.get("completionTime")


# job_test.go:691-810 TestBackoffForOnFailure ------------------------------------------------
def test_backoff_for_on_failure():
    job = new_tfjob(4, 2, backoffLimit=4)
    for spec in job["spec"]["tfReplicaSpecs"].values():
        spec["restartPolicy"] = c.RESTART_ON_FAILURE
    ctrl, store = make_controller(TFJobReconciler, job)
    set_pods_statuses(store, job, "worker", 0, 4, 0, 0, restart_counts=[1, 1, 1, 1])
    set_pods_statuses(store, job, "ps", 0, 2, 0, 0, restart_counts=[0, 0])
    set_services(store, job, "worker", 4)
    set_services(store, job, "ps", 2)
    status = _reconcile(ctrl, job)
    assert len(ctrl.pod_control.delete_pod_name) == 6
    assert len(ctrl.service_control.delete_service_name) == 6
    failed = c.get_condition(status, c.JOB_FAILED)
    assert failed and "reached the specified backoff limit" in failed["message"]
    # below the limit nothing is torn down
    job2 = new_tfjob(4, 2, backoffLimit=5, name="below")
    for spec in job2["spec"]["tfReplicaSpecs"].values():
        spec["restartPolicy"] = c.RESTART_ON_FAILURE
    ctrl2, store2 = make_controller(TFJobReconciler, job2)
    set_pods_statuses(store2, job2, "worker", 0, 4, 0, 0, restart_counts=[1, 1, 1, 1])
    set_pods_statuses(store2, job2, "ps", 0, 2, 0, 0)
    st2 = _reconcile(ctrl2, job2)
    assert not ctrl2.pod_control.delete_pod_name and not c.is_failed(st2)


# pod_test.go:109-228 TestExpectation / TestExpectationWithError ------------------------------------
def test_expectation_raised_per_replica_type_and_rolled_back_on_error():
    job = new_tfjob(2, 1)
    ctrl, store = make_controller(TFJobReconciler, job)
    _reconcile(ctrl, job)
    key = c.key_of(job)
    assert ctrl.expectations.get(exp.gen_expectation_pods_key(key, "worker")) == (2, 0)
    assert ctrl.expectations.get(exp.gen_expectation_pods_key(key, "ps")) == (1, 0)
    assert ctrl.expectations.get(exp.gen_expectation_services_key(key, "worker")) == (2, 0)
    assert not exp.satisfied_expectations(ctrl.expectations, key, ["Worker", "PS"])
    for _ in range(2):
        ctrl.expectations.creation_observed(exp.gen_expectation_pods_key(key, "worker"))
    assert exp.satisfied_expectations(ctrl.expectations, key, ["Worker"])
    # creation failure: the expectation is decremented again
    job2 = new_tfjob(1, 0, name="err")
    ctrl2, _ = make_controller(TFJobReconciler, job2)
    ctrl2.pod_control.err = RuntimeError("fake create error")
    with pytest.raises(RuntimeError):
        _reconcile(ctrl2, job2)
    assert ctrl2.expectations.get(exp.gen_expectation_pods_key(c.key_of(job2), "worker")) == (0, 0)


# pod_test.go:230-351 TestClusterSpec ------------------------------------------------
@pytest.mark.parametrize("worker,ps,evaluator,ns,domain,expected", [
    (1, 0, 0, "ns0", "", ""),
    (1, 0, 0, "ns1", "tf.training.com", ""),
    (1, 1, 0, "ns2", "tf.training.org",
     '{"cluster":{"ps":["%(n)s-ps-0.ns2.svc.tf.training.org:2222"],"worker":["%(n)s-worker-0.ns2.svc.tf.training.org:2222"]},"task":{"type":"worker","index":0},"environment":"cloud"}'),
    (1, 1, 1, "ns3", "tf.training.io",
     '{"cluster":{"evaluator":["%(n)s-evaluator-0.ns3.svc.tf.training.io:2222"],"ps":["%(n)s-ps-0.ns3.svc.tf.training.io:2222"],"worker":["%(n)s-worker-0.ns3.svc.tf.training.io:2222"]},"task":{"type":"worker","index":0},"environment":"cloud"}'),
    (1, 1, 1, "ns3", "",
     '{"cluster":{"evaluator":["%(n)s-evaluator-0.ns3.svc:2222"],"ps":["%(n)s-ps-0.ns3.svc:2222"],"worker":["%(n)s-worker-0.ns3.svc:2222"]},"task":{"type":"worker","index":0},"environment":"cloud"}'),
])
def test_cluster_spec(monkeypatch, worker, ps, evaluator, ns, domain, expected):
    monkeypatch.setenv(c.ENV_CUSTOM_CLUSTER_DOMAIN, domain)
    job = tfapi.set_defaults(new_tfjob(worker, ps, evaluator=evaluator, namespace=ns))
    ctrl, _ = make_controller(TFJobReconciler)
    tmpl = copy.deepcopy(job["spec"]["tfReplicaSpecs"]["Worker"]["template"])
    ctrl.set_cluster_spec(job, tmpl, "worker", "0")
    env = tmpl["spec"]["containers"][0].get("env") or []
    if expected == "":
        assert env == []
    else:
        assert env[0]["name"] == "TF_CONFIG"
        assert env[0]["value"] == expected % {"n": TEST_JOB_NAME}


def test_sparse_cluster_spec_for_dynamic_workers():
    """tensorflow.go:64-83,110-119 (enableDynamicWorker)."""
    job = tfapi.set_defaults(new_tfjob(3, 2, enableDynamicWorker=True))
    n = TEST_JOB_NAME
    assert gen_tf_config_json(job, "worker", "1") == (
        '{"sparseCluster":{"worker":{"1":"%s-worker-1.default.svc:2222"},"ps":["%s-ps-0.default.svc:2222",'
        '"%s-ps-1.default.svc:2222"]},"task":{"type":"worker","index":1}}' % (n, n, n))
    assert gen_tf_config_json(job, "ps", "1") == (
        '{"sparseCluster":{"worker":{},"ps":["%s-ps-1.default.svc:2222"]},"task":{"type":"ps","index":1}}' % n)


# pod_test.go:353-382 TestIsDistributed ------------------------------------------------
@pytest.mark.parametrize("kw,expected", [
    (dict(worker=1, ps=0), False), (dict(worker=1, ps=1), True), (dict(worker=0, ps=0, chief=1), False),
    (dict(worker=1, ps=0, chief=1), True), (dict(worker=0, ps=0, master=1), False),
    (dict(worker=2, ps=0), True), (dict(worker=0, ps=0, evaluator=1), False),
])
def test_is_distributed(kw, expected):
    assert is_distributed(new_tfjob(**kw)) is expected


# pod_test.go:384-440 TestRestartPolicy ------------------------------------------------
@pytest.mark.parametrize("policy,expected", [
    (c.RESTART_EXIT_CODE, "Never"), (c.RESTART_NEVER, "Never"), (c.RESTART_ALWAYS, "Always"),
    (c.RESTART_ON_FAILURE, "OnFailure")])
def test_restart_policy(policy, expected):
    job = new_tfjob(1, 0)
    job["spec"]["tfReplicaSpecs"]["Worker"]["restartPolicy"] = policy
    ctrl, _ = make_controller(TFJobReconciler, job)
    _reconcile(ctrl, job)
    assert ctrl.pod_control.templates[0]["spec"]["restartPolicy"] == expected


def test_template_restart_policy_is_overwritten_with_warning():
    job = new_tfjob(1, 0)
    job["spec"]["tfReplicaSpecs"]["Worker"]["template"]["spec"]["restartPolicy"] = "Always"
    ctrl, _ = make_controller(TFJobReconciler, job)
    _reconcile(ctrl, job)
    assert ctrl.pod_control.templates[0]["spec"]["restartPolicy"] == "Never"
    assert any("SettedPodTemplateRestartPolicy" in e for e in ctrl.recorder.events)


# pod_test.go:442-527 TestExitCode ------------------------------------------------
def test_exit_code_retryable_pod_is_deleted_and_job_restarting():
    job = new_tfjob(1, 0)
    job["spec"]["tfReplicaSpecs"]["Worker"]["restartPolicy"] = c.RESTART_EXIT_CODE
    ctrl, store = make_controller(TFJobReconciler, job)
    store.create("pods", new_pod(job, "worker", 0, c.POD_FAILED, exit_code=130))
    status = _reconcile(ctrl, job)
    assert ctrl.pod_control.delete_pod_name == [f"{TEST_JOB_NAME}-worker-0"]
    assert c.has_condition(status, c.JOB_RESTARTING) and not c.is_failed(status)
    assert c.get_condition(status, c.JOB_RESTARTING)["message"] == \
        f"TFJob {TEST_JOB_NAME} is restarting because Worker replica(s) failed."
    assert any("ExitedWithCode" in e for e in ctrl.recorder.events)


def test_exit_code_permanent_pod_fails_the_job():
    job = new_tfjob(1, 0)
    job["spec"]["tfReplicaSpecs"]["Worker"]["restartPolicy"] = c.RESTART_EXIT_CODE
    ctrl, store = make_controller(TFJobReconciler, job)
    store.create("pods", new_pod(job, "worker", 0, c.POD_FAILED, exit_code=1))
    status = _reconcile(ctrl, job)
    assert ctrl.pod_control.delete_pod_name == []
    assert c.is_failed(status)
    assert c.get_condition(status, c.JOB_FAILED)["message"] == \
        f"TFJob default/{TEST_JOB_NAME} has failed because 1 Worker replica(s) failed."


@pytest.mark.parametrize("code,retry", [(1, False), (2, False), (126, False), (127, False), (128, False),
                                        (139, False), (130, True), (137, True), (143, True), (138, True),
                                        (0, False), (3, False), (255, False)])
def test_is_retryable_exit_code(code, retry):
    assert c.is_retryable_exit_code(code) is retry


# pod_test.go:530-685 TestScaleDown / TestScaleUp ------------------------------------------------
def test_scale_down_deletes_out_of_range_pods():
    job = new_tfjob(2, 0, enableDynamicWorker=True)
    ctrl, store = make_controller(TFJobReconciler, job)
    for i in range(3):
        store.create("pods", new_pod(job, "worker", i, c.POD_RUNNING))
    _reconcile(ctrl, job)
    assert ctrl.pod_control.delete_pod_name == [f"{TEST_JOB_NAME}-worker-2"]
    assert ctrl.pod_control.templates == []


def test_scale_up_creates_missing_indices():
    job = new_tfjob(3, 0, enableDynamicWorker=True)
    ctrl, store = make_controller(TFJobReconciler, job)
    store.create("pods", new_pod(job, "worker", 0, c.POD_RUNNING))
    _reconcile(ctrl, job)
    names = [t["metadata"]["name"] for t in ctrl.pod_control.templates]
    assert names == [f"{TEST_JOB_NAME}-worker-1", f"{TEST_JOB_NAME}-worker-2"]
    # dynamic workers get the sparse TF_CONFIG
    assert '"sparseCluster"' in ctrl.pod_control.templates[0]["spec"]["containers"][0]["env"][0]["value"]


# pod_test.go:687-823 TestIsWorker0Completed ------------------------------------------------
@pytest.mark.parametrize("workers,codes,phases,expected", [
    (1, {0: 0}, {0: c.POD_SUCCEEDED}, True),
    (1, {0: 1}, {0: c.POD_FAILED}, False),
    (2, {0: 0}, {0: c.POD_SUCCEEDED, 1: c.POD_RUNNING}, True),
    (2, {1: 0}, {0: c.POD_RUNNING, 1: c.POD_SUCCEEDED}, False),
    (2, {}, {0: c.POD_SUCCEEDED, 1: c.POD_SUCCEEDED}, False),   # succeeded without an exit code
])
def test_is_worker0_completed(workers, codes, phases, expected):
    job = new_tfjob(workers, 1)
    ctrl, store = make_controller(TFJobReconciler, job)
    for i, ph in phases.items():
        store.create("pods", new_pod(job, "worker", i, ph, exit_code=codes.get(i)))
    tfapi.set_defaults(job)
    assert ctrl.is_worker0_completed(job, c.replica_specs(job, ctrl.info)) is expected


def test_no_worker_spec_counts_as_worker0_completed():
    job = tfapi.set_defaults(new_tfjob(0, 0, chief=1))
    ctrl, _ = make_controller(TFJobReconciler, job)
    assert ctrl.is_worker0_completed(job, c.replica_specs(job, ctrl.info)) is True


# status_test.go:97-508 TestStatus (19 rows) ------------------------------------------------
STATUS_ROWS = [
    ("Chief worker is succeeded", dict(worker=1, ps=0, chief=1), {"SW": 1, "SC": 1}, False, False, c.JOB_SUCCEEDED),
    ("Chief worker is running", dict(worker=1, ps=0, chief=1), {"AC": 1}, False, False, c.JOB_RUNNING),
    ("Chief worker is failed", dict(worker=1, ps=0, chief=1), {"FC": 1}, False, False, c.JOB_FAILED),
    ("(No chief worker) Worker is failed", dict(worker=1, ps=0), {"FW": 1}, False, False, c.JOB_FAILED),
    ("(No chief worker) Worker is succeeded", dict(worker=1, ps=0), {"SW": 1}, False, False, c.JOB_SUCCEEDED),
    ("(No chief worker) Worker is running", dict(worker=1, ps=0), {"AW": 1}, False, False, c.JOB_RUNNING),
    ("2 workers succeeded, 2 active", dict(worker=4, ps=2), {"AP": 2, "SW": 2, "AW": 2}, False, False, c.JOB_RUNNING),
    ("2 workers running, 2 failed", dict(worker=4, ps=2), {"AP": 2, "FW": 2, "AW": 2}, False, False, c.JOB_FAILED),
    ("2 workers succeeded, 2 failed", dict(worker=4, ps=2), {"AP": 2, "FW": 2, "SW": 2}, False, False, c.JOB_FAILED),
    ("worker-0 succeeded, 3 active", dict(worker=4, ps=2), {"AP": 2, "SW": 1, "AW": 3}, False, True, c.JOB_SUCCEEDED),
    ("AllWorkers: worker-0 succeeded, 3 active", dict(worker=4, ps=0, successPolicy="AllWorkers"), {"SW": 1, "AW": 3}, False, True, c.JOB_RUNNING),
    ("AllWorkers: 4 succeeded", dict(worker=4, ps=0, successPolicy="AllWorkers"), {"SW": 4}, False, True, c.JOB_SUCCEEDED),
    ("AllWorkers: worker-0 succeeded, 2 running, 1 failed", dict(worker=4, ps=0, successPolicy="AllWorkers"), {"FW": 1, "SW": 1, "AW": 2}, False, True, c.JOB_FAILED),
    ("Chief running, workers failed", dict(worker=4, ps=2, chief=1), {"AP": 2, "FW": 4, "AC": 1}, False, False, c.JOB_RUNNING),
    ("Chief running, workers succeeded", dict(worker=4, ps=2, chief=1), {"AP": 2, "SW": 4, "AC": 1}, False, False, c.JOB_RUNNING),
    ("Chief running, a PS failed", dict(worker=4, ps=2, chief=1), {"FP": 1, "AP": 1, "SW": 4, "AC": 1}, False, False, c.JOB_FAILED),
    ("Chief failed, workers succeeded", dict(worker=4, ps=2, chief=1), {"AP": 2, "SW": 4, "FC": 1}, False, False, c.JOB_FAILED),
    ("Chief succeeded, workers failed", dict(worker=4, ps=2, chief=1), {"AP": 2, "FW": 4, "SC": 1}, False, False, c.JOB_SUCCEEDED),
    ("Chief failed and restarting", dict(worker=4, ps=2, chief=1), {"AP": 2, "FW": 4, "FC": 1}, True, False, c.JOB_RESTARTING),
]


@pytest.mark.parametrize("row", STATUS_ROWS, ids=[r[0] for r in STATUS_ROWS])
def test_status(row):
    _, kw, counts, restart, worker0_completed, expected = row
    job = new_tfjob(**kw)
    if restart:
        for spec in job["spec"]["tfReplicaSpecs"].values():
            spec["restartPolicy"] = c.RESTART_EXIT_CODE
    ctrl, store = make_controller(TFJobReconciler, job)
    for letter, typ in (("P", "ps"), ("W", "worker"), ("C", "chief")):
        idx = 0
        for _ in range(counts.get("S" + letter, 0)):      # succeeded first, like setStatusForTest
            code = 0 if (worker0_completed and typ == "worker" and idx == 0) else None
            store.create("pods", new_pod(job, typ, idx, c.POD_SUCCEEDED, exit_code=code))
            idx += 1
        for _ in range(counts.get("F" + letter, 0)):
            store.create("pods", new_pod(job, typ, idx, c.POD_FAILED, exit_code=130 if restart else None))
            idx += 1
        for _ in range(counts.get("A" + letter, 0)):
            store.create("pods", new_pod(job, typ, idx, c.POD_RUNNING))
            idx += 1
    status = _reconcile(ctrl, job)
    assert c.get_condition(status, expected) is not None, [cd["type"] for cd in status["conditions"]]
    # filterOutConditionTest (status_test.go:585-592)
    if c.is_failed(status) or c.is_succeeded(status):
        assert not any(cd["type"] == c.JOB_RUNNING and cd["status"] == "True"
                       for cd in status["conditions"])


def test_condition_list_algebra():
    st = c.new_status()
    c.update_job_conditions(st, c.JOB_CREATED, "r", "m")
    c.update_job_conditions(st, c.JOB_RUNNING, "r", "m")
    c.update_job_conditions(st, c.JOB_RESTARTING, "r", "m")
    assert [x["type"] for x in st["conditions"]] == [c.JOB_CREATED, c.JOB_RESTARTING]
    c.update_job_conditions(st, c.JOB_RUNNING, "r", "m")
    assert [x["type"] for x in st["conditions"]] == [c.JOB_CREATED, c.JOB_RUNNING]
    t0 = c.get_condition(st, c.JOB_RUNNING)["lastTransitionTime"]
    c.update_job_conditions(st, c.JOB_RUNNING, "r", "m2")          # identical type/status/reason: no-op
    assert c.get_condition(st, c.JOB_RUNNING)["message"] == "m"
    c.update_job_conditions(st, c.JOB_RUNNING, "r2", "m3", now="2030-01-01T00:00:00Z")
    assert c.get_condition(st, c.JOB_RUNNING)["lastTransitionTime"] == t0   # status unchanged
    c.update_job_conditions(st, c.JOB_SUCCEEDED, "r", "m")
    assert st["conditions"][-1]["type"] == c.JOB_SUCCEEDED
    assert c.get_condition(st, c.JOB_RUNNING)["status"] == "False"
    c.update_job_conditions(st, c.JOB_FAILED, "r", "m")
    assert st["conditions"][-1]["type"] == c.JOB_FAILED
    before = copy.deepcopy(st)
    c.update_job_conditions(st, c.JOB_RUNNING, "r", "m")            # Failed is sticky
    assert st == before


# pod.go:168-258 createNewPod: labels, names, gang annotations --------------------------------------
def test_created_pod_labels_names_and_master_role():
    job = new_tfjob(2, 1)
    ctrl, _ = make_controller(TFJobReconciler, job)
    _reconcile(ctrl, job)
    by_name = {t["metadata"]["name"]: t for t in ctrl.pod_control.templates}
    assert set(by_name) == {f"{TEST_JOB_NAME}-worker-0", f"{TEST_JOB_NAME}-worker-1", f"{TEST_JOB_NAME}-ps-0"}
    w0 = by_name[f"{TEST_JOB_NAME}-worker-0"]["metadata"]["labels"]
    assert w0 == {"group-name": "kubeflow.org", "job-name": TEST_JOB_NAME, "replica-type": "worker",
                  "replica-index": "0", "job-role": "master"}     # worker-0 is master without a chief
    assert "job-role" not in by_name[f"{TEST_JOB_NAME}-worker-1"]["metadata"]["labels"]
    svc = {s["metadata"]["name"]: s for s in ctrl.service_control.services}
    s0 = svc[f"{TEST_JOB_NAME}-ps-0"]
    assert s0["spec"]["clusterIP"] == "None"
    assert s0["spec"]["selector"] == s0["metadata"]["labels"]
    assert s0["spec"]["ports"] == [{"name": "tfjob-port", "port": 2222}]


def test_gang_scheduling_annotations_and_podgroup():
    job = new_tfjob(2, 1)
    job["spec"]["runPolicy"] = {"schedulingPolicy": {"minAvailable": 2, "queue": "q1",
                                                      "priorityClass": "high"}}
    ctrl, store = make_controller(TFJobReconciler, job, gang=True)
    _reconcile(ctrl, job)
    t = ctrl.pod_control.templates[0]
    assert t["spec"]["schedulerName"] == "volcano"
    assert t["metadata"]["annotations"] == {"scheduling.k8s.io/group-name": TEST_JOB_NAME,
                                            "volcano.sh/task-spec": t["metadata"]["labels"]["replica-type"]}
    pg = store.get("podgroups", "default", TEST_JOB_NAME)
    assert pg["spec"] == {"minMember": 2, "queue": "q1", "priorityClassName": "high"}
    assert pg["metadata"]["ownerReferences"][0]["name"] == TEST_JOB_NAME
    # without minAvailable the group needs every replica
    job2 = new_tfjob(4, 2, name="all")
    ctrl2, store2 = make_controller(TFJobReconciler, job2, gang=True)
    _reconcile(ctrl2, job2)
    assert store2.get("podgroups", "default", "all")["spec"]["minMember"] == 6


def test_adoption_and_release():
    """ClaimPods (tfjob_controller.go:251-289): orphans matching the selector are adopted,
    owned pods whose labels drifted are released."""
    job = new_tfjob(1, 0)
    ctrl, store = make_controller(TFJobReconciler, job)
    orphan = new_pod(job, "worker", 0, c.POD_RUNNING)
    orphan["metadata"]["ownerReferences"] = []
    store.create("pods", orphan)
    drift = new_pod(job, "worker", 7, c.POD_RUNNING)
    drift["metadata"]["labels"]["job-name"] = "someone-else"
    store.create("pods", drift)
    pods = ctrl.get_pods_for_job(job)
    assert [c.name_of(p) for p in pods] == [f"{TEST_JOB_NAME}-worker-0"]
    assert store.get("pods", "default", f"{TEST_JOB_NAME}-worker-0")["metadata"]["ownerReferences"][0]["uid"] \
        == job["metadata"]["uid"]
    assert not store.get("pods", "default", f"{TEST_JOB_NAME}-worker-7")["metadata"].get("ownerReferences")
