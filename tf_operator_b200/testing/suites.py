"""The e2e suites of the reference (py/kubeflow/tf_operator/*_tests.py) as
reflection-runner test classes against the single-box cluster. Each class mirrors one reference
module; the runner instantiates it with the parsed args and the cluster and calls every
``test_*`` method.
"""
from __future__ import annotations

import json
import logging
import time

from ..api import common as c
from ..store.memory import ApiError
from . import fixtures, k8s_util, tf_job_client, util


class E2ETestCase:
    """Base of all suites (reference: test_util.TestCase + the per-module ctor that stores
    args.namespace / args.params / args.app_dir)."""

    def __init__(self, args, cluster):
        self.args = args
        self.cluster = cluster
        self.namespace = getattr(args, "namespace", None) or "default"
        self.name = None
        self.tfjob_version = getattr(args, "tfjob_version", "v1")

    # helpers shared by the suites -------------------------------------------------------------
    def create(self, component: str, name: str = None) -> dict:
        self.name = name or component.replace("_", "-")
        job = fixtures.build(component, self.name, self.namespace)
        return tf_job_client.create_tf_job(self.cluster, job, self.tfjob_version)

    def wait_all_running(self, n: int) -> None:
        sel = tf_job_client.to_selector(tf_job_client.get_labels(self.name))
        deadline = time.time() + 60
        while time.time() < deadline:
            pods = k8s_util.list_pods(self.cluster.store, self.namespace, sel)
            if len(pods) == n and all(p["status"].get("phase") == "Running" for p in pods):
                break
            time.sleep(0.02)
        else:
            raise util.TimeoutError_(f"{self.name}: {n} running pods expected")
        for p in k8s_util.list_pods(self.cluster.store, self.namespace, sel):
            util.send_request(self.cluster.api.url, self.namespace, c.name_of(p), "", {})

    def delete_and_wait(self) -> None:
        tf_job_client.delete_tf_job(self.cluster, self.namespace, self.name, self.tfjob_version)
        tf_job_client.wait_for_delete(self.cluster, self.namespace, self.name, self.tfjob_version)
        k8s_util.wait_for_pods_to_be_deleted(
            self.cluster.store, self.namespace,
            tf_job_client.to_selector(tf_job_client.get_labels(self.name)))

    def finished(self) -> dict:
        job = tf_job_client.wait_for_job(self.cluster, self.namespace, self.name, self.tfjob_version,
                                         status_callback=tf_job_client.log_status)
        logging.info("Final TFJob:\n%s", json.dumps(job.get("status", {}), default=str)[:2000])
        return job


class SimpleTfJobTests(E2ETestCase):
    """simple_tfjob_tests.py:26-90: run to success, check the creation events, delete, and do it
    again under the same name (num_trials)."""

    def run_simple_tfjob(self, component: str):
        for trial in range(getattr(self.args, "num_trials", 2) or 2):
            logging.info("Trial %s", trial)
            self.create(component)
            n = sum(s["replicas"] for s in fixtures.build(component)["spec"]["tfReplicaSpecs"].values())
            self.wait_all_running(n)
            chief = "chief" if "master_is_chief" in component or "simple" in component else "worker"
            tf_job_client.terminate_replicas(self.cluster, self.namespace, self.name, chief, 1)
            job = self.finished()
            assert tf_job_client.job_succeeded(job), f"job {self.name} did not succeed"
            failures = tf_job_client.get_creation_failures_from_tfjob(self.cluster, self.namespace, job)
            assert not failures, failures
            self.delete_and_wait()

    def test_simple_tfjob_cpu(self):
        self.run_simple_tfjob("simple_tfjob_v1")


class ShutdownPolicyTests(E2ETestCase):
    """shutdown_policy_tests.py:25-93: the job ends when the chief (or worker 0 when there is
    no chief) exits, whatever the other replicas do."""

    def run_tfjob_with_shutdown_policy(self, component: str, replica: str):
        self.create(component)
        n = sum(s["replicas"] for s in fixtures.build(component)["spec"]["tfReplicaSpecs"].values())
        self.wait_all_running(n)
        tf_job_client.terminate_replicas(self.cluster, self.namespace, self.name, replica, 1)
        job = self.finished()
        assert tf_job_client.job_succeeded(job), f"job {self.name} did not succeed"
        self.delete_and_wait()

    def test_shutdown_chief(self):
        self.run_tfjob_with_shutdown_policy("master_is_chief_v1", "chief")

    def test_shutdown_worker0(self):
        self.run_tfjob_with_shutdown_policy("worker0_is_chief_v1", "worker")


class CleanPodPolicyTests(E2ETestCase):
    """cleanpod_policy_tests.py:26-115: after the job finishes, All deletes every pod, Running
    deletes the still-running PS pods, None keeps all seven."""

    def run_tfjob_with_cleanpod_policy(self, component: str, policy: str):
        self.create(component)
        job = self.finished()
        assert tf_job_client.job_succeeded(job)
        sel = tf_job_client.to_selector(tf_job_client.get_labels(self.name))
        store = self.cluster.store

        def names():
            return {c.name_of(p) for p in k8s_util.list_pods(store, self.namespace, sel)}
        if policy == "All":
            k8s_util.wait_for_pods_to_be_deleted(store, self.namespace, sel)
        elif policy == "Running":
            k8s_util.wait_for_pods_to_be_deleted(
                store, self.namespace, tf_job_client.to_selector(tf_job_client.get_labels(self.name, "ps")))
            k8s_util.wait_for_pods_to_be_in_phases(store, self.namespace, sel, ["Succeeded"])
            assert len(names()) == 5, names()       # chief + 4 workers stay
        else:
            time.sleep(0.5)
            assert len(names()) == 7, names()
            k8s_util.wait_for_pods_to_be_in_phases(
                store, self.namespace,
                tf_job_client.to_selector(tf_job_client.get_labels(self.name, "ps")), ["Running"])
        self.delete_and_wait()

    def test_cleanpod_all(self):
        self.run_tfjob_with_cleanpod_policy("clean_pod_all_v1", "All")

    def test_cleanpod_running(self):
        self.run_tfjob_with_cleanpod_policy("clean_pod_running_v1", "Running")

    def test_cleanpod_none(self):
        self.run_tfjob_with_cleanpod_policy("clean_pod_none_v1", "None")


class ReplicaRestartPolicyTests(E2ETestCase):
    """replica_restart_policy_tests.py:23-120: kill ps-0 with a given exit code and check
    whether its container is started again, per restartPolicy."""

    def run_tfjob_with_replica_restart_policy(self, component, exit_code, expect_restart, final):
        self.create(component)
        self.wait_all_running(3)
        ok = tf_job_client.terminate_and_verify_start_time(
            self.cluster, self.namespace, self.name, "ps", 0, exit_code, expect_restart)
        assert ok, f"{component}: exit {exit_code}, restart expected={expect_restart}"
        if final:
            job = self.finished()
            assert job["status"]["conditions"][-1]["type"] == final, job["status"]["conditions"]
        self.delete_and_wait()

    def test_restart_always_exit_code_0(self):
        self.run_tfjob_with_replica_restart_policy("replica_restart_policy_always_v1", 0, True, None)

    def test_restart_onfailure_exit_code_1(self):
        self.run_tfjob_with_replica_restart_policy("replica_restart_policy_onfailure_v1", 1, True, None)

    def test_restart_never_exit_code_1(self):
        self.run_tfjob_with_replica_restart_policy("replica_restart_policy_never_v1", 1, False, "Failed")

    def test_restart_exitcode_permanent(self):
        # 1 is a permanent error under ExitCode: no restart, job fails
        self.run_tfjob_with_replica_restart_policy("replica_restart_policy_exitcode_v1", 1, False, "Failed")

    def test_restart_exitcode_retryable(self):
        # 130 (SIGINT) is retryable under ExitCode: the pod is re-created
        self.run_tfjob_with_replica_restart_policy("replica_restart_policy_exitcode_v1", 130, True, None)


class EstimatorRunconfigTests(E2ETestCase):
    """estimator_runconfig_tests.py:15-130: every replica's RunConfig, as parsed from the
    TF_CONFIG the operator injected, has the right task, master, counts and cluster spec."""

    def test_tfjob_and_verify_runconfig(self):
        self.create("estimator_runconfig_v1")
        self.wait_all_running(6)
        ns, name = self.namespace, self.name

        def host(rt, i):
            return f"{name}-{rt}-{i}.{ns}.svc:2222"
        cluster_spec = {"chief": [host("chief", 0)], "ps": [host("ps", 0), host("ps", 1)],
                        "worker": [host("worker", 0), host("worker", 1)]}
        for rt, n in (("chief", 1), ("ps", 2), ("worker", 2), ("evaluator", 1)):
            for i in range(n):
                raw = util.send_request(self.cluster.api.url, ns, f"{name}-{rt}-{i}", "runconfig", {})
                got = json.loads(raw)
                assert got["task_type"] == rt and got["task_id"] == i, got
                assert got["is_chief"] == (rt == "chief"), got
                if rt == "evaluator":
                    assert got["master"] == "" and got["cluster_spec"] == {}, got
                else:
                    assert got["master"] == "grpc://" + host(rt, i), got
                    assert got["cluster_spec"] == cluster_spec, got
                    assert got["num_ps_replicas"] == 2, got
                    assert got["num_worker_replicas"] == 3, got     # chief counts as a worker
        tf_job_client.terminate_replicas(self.cluster, ns, name, "chief", 1)
        assert tf_job_client.job_succeeded(self.finished())
        self.delete_and_wait()


class InvalidTfJobTests(E2ETestCase):
    """invalid_tfjob_tests.py:17-80. With the all-in-one operator's CRD schema the API server
    rejects the object; the legacy (schema-less) path instead marks the job Failed with reason
    InvalidTFJobSpec - tests/test_e2e_launcher.py covers that flavour."""

    def test_invalid_tfjob_spec(self):
        try:
            self.create("invalid_tfjob_v1")
        except ApiError as e:
            assert e.code in (400, 422), e
            return
        job = self.finished()
        last = job["status"]["conditions"][-1]
        assert last["type"] == "Failed" and "invalid" in last.get("reason", "").lower(), last
        self.delete_and_wait()


class PodNamesValidationTests(E2ETestCase):
    """pod_names_validation_tests.py:14,47-84: pods and services are named
    {job}-{replica type}-{index}."""

    def test_pod_names(self):
        self.create("pod_names_validation_v1")
        self.wait_all_running(7)
        expected = {f"{self.name}-chief-0", f"{self.name}-ps-0", f"{self.name}-ps-1",
                    f"{self.name}-worker-0", f"{self.name}-worker-1", f"{self.name}-worker-2",
                    f"{self.name}-evaluator-0"}
        actual = tf_job_client.get_pod_names(self.cluster, self.namespace, self.name)
        assert actual == expected, f"Actual pod names doesn't match. Expected: {expected} Actual: {actual}"
        tf_job_client.terminate_replicas(self.cluster, self.namespace, self.name, "chief", 1)
        self.finished()
        self.delete_and_wait()


class DistributedTrainingJobTests(E2ETestCase):
    """distributed_training_tests.py:14-70: three workers train with a collective all-reduce
    and the job succeeds."""

    def test_distributed_training_independent_worker(self):
        self.create("distributed_training_v1")
        job = self.finished()
        assert tf_job_client.job_succeeded(job), job["status"]
        failures = tf_job_client.get_creation_failures_from_tfjob(self.cluster, self.namespace, job)
        assert not failures, failures
        self.delete_and_wait()


SUITES = {
    "simple_tfjob_tests": SimpleTfJobTests,
    "shutdown_policy_tests": ShutdownPolicyTests,
    "cleanpod_policy_tests": CleanPodPolicyTests,
    "replica_restart_policy_tests": ReplicaRestartPolicyTests,
    "estimator_runconfig_tests": EstimatorRunconfigTests,
    "invalid_tfjob_tests": InvalidTfJobTests,
    "pod_names_validation_tests": PodNamesValidationTests,
    "distributed_training_tests": DistributedTrainingJobTests,
}
