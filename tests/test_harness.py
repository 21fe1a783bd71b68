"""The e2e harness package (tf_operator_b200/testing: SURVEY.md E1-E5, E8): fixtures, helper
client, junit writer and the reflection runner, exercised against the single-box cluster."""
import os
import subprocess
import sys
from xml.etree import ElementTree as ET

import pytest

from tf_operator_b200 import api
from tf_operator_b200.testing import fixtures, k8s_util, test_runner, tf_job_client, util
from tf_operator_b200.testing.suites import SUITES

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_fifteen_fixtures_match_the_reference_components():
    assert sorted(fixtures.COMPONENTS) == sorted([
        "simple_tfjob_v1", "gpu_tfjob_v1", "master_is_chief_v1", "worker0_is_chief_v1",
        "clean_pod_all_v1", "clean_pod_running_v1", "clean_pod_none_v1",
        "replica_restart_policy_always_v1", "replica_restart_policy_onfailure_v1",
        "replica_restart_policy_never_v1", "replica_restart_policy_exitcode_v1",
        "estimator_runconfig_v1", "distributed_training_v1", "invalid_tfjob_v1",
        "pod_names_validation_v1"])
    simple = fixtures.build("simple_tfjob_v1", "s", "ns1")
    assert simple["metadata"] == {"name": "s", "namespace": "ns1"}
    assert {k: v["replicas"] for k, v in simple["spec"]["tfReplicaSpecs"].items()} == \
        {"Chief": 1, "PS": 2, "Worker": 4}
    gpu = fixtures.build("gpu_tfjob_v1")
    assert gpu["spec"]["tfReplicaSpecs"]["Chief"]["template"]["spec"]["containers"][0]["resources"] == \
        {"limits": {"nvidia.com/gpu": 1}}
    assert fixtures.build("clean_pod_running_v1")["spec"]["cleanPodPolicy"] == "Running"
    for name in fixtures.COMPONENTS:
        job = fixtures.build(name)
        if name == "invalid_tfjob_v1":
            with pytest.raises(Exception):
                api.validate(api.set_defaults(job) or job)
        else:
            api.set_defaults(job)
            api.validate(job)


def test_labels_selector_and_event_parsing():
    assert tf_job_client.get_labels("j", "Worker", 2) == {
        "group-name": "kubeflow.org", "job-name": "j", "replica-type": "worker", "replica-index": "2"}
    assert tf_job_client.to_selector({"a": "1", "b": "2"}) == "a=1,b=2"
    events = [{"message": "Created pod: j-worker-0"}, {"message": "Created service: j-worker-0"},
              {"message": "Created pod: j-ps-0"}, {"message": "Deleted pod: j-ps-0"}]
    pods, svcs = k8s_util.parse_events(events)
    assert pods == {"j-worker-0", "j-ps-0"} and svcs == {"j-worker-0"}
    assert tf_job_client.job_succeeded({"status": {"conditions": [{"type": "Running"}, {"type": "Succeeded"}]}})
    assert not tf_job_client.job_succeeded({"status": {"conditions": [{"type": "Succeeded"}, {"type": "Failed"}]}})
    assert not tf_job_client.job_succeeded({"status": {}})


def test_junit_writer_counts_failures_and_unrun_cases(tmp_path):
    ok, bad, unrun = util.TestCase("C", "test_ok"), util.TestCase("C", "test_bad"), util.TestCase("C", "test_x")
    util.wrap_test(lambda: None, ok)
    with pytest.raises(ValueError):
        util.wrap_test(lambda: (_ for _ in ()).throw(ValueError("boom")), bad)
    assert ok.failure is None and ok.time is not None
    assert "ValueError: boom" in bad.failure
    path = tmp_path / "junit.xml"
    util.create_junit_xml_file([ok, bad, unrun], str(path))
    root = ET.parse(path).getroot()
    assert root.attrib["tests"] == "3" and root.attrib["failures"] == "1"
    fails = {tc.attrib["name"]: tc.find("failure") is not None for tc in root.findall("testcase")}
    assert fails == {"test_ok": False, "test_bad": True, "test_x": True}   # never run => failed
    assert util.get_num_failures(path.read_text()) == 1
    suite = util.TestSuite("C")
    suite.create("a")
    with pytest.raises(ValueError):
        suite.create("a")
    with pytest.raises(ValueError):
        suite.get("missing")


def test_runner_retries_flakes_but_not_assertions():
    class Args:
        num_trials = 2
    calls = {"n": 0}

    def flaky():
        calls["n"] += 1
        if calls["n"] == 1:
            raise RuntimeError("apiserver hiccup")
    tc = util.TestCase("C", "test_flaky")
    test_runner.run_test(tc, flaky, Args)
    assert tc.failure is None and calls["n"] == 3           # 1 failed attempt + 2 trials

    def wrong():
        calls["n"] += 1
        raise AssertionError("pods differ")
    calls["n"] = 0
    tc2 = util.TestCase("C", "test_wrong")
    test_runner.run_test(tc2, wrong, Args)
    assert "pods differ" in tc2.failure and calls["n"] == 1  # a verdict is not retried
    assert test_runner._test_methods(SUITES["shutdown_policy_tests"]) == \
        ["test_shutdown_chief", "test_shutdown_worker0"]


def test_runner_cli_runs_two_suites_and_writes_junit(tmp_path):
    res = subprocess.run([sys.executable, "-m", "tf_operator_b200.testing.test_runner", "--suite",
                          "shutdown_policy_tests,pod_names_validation_tests", "--artifacts_path",
                          str(tmp_path)], cwd=REPO, capture_output=True, text=True, timeout=300)
    assert res.returncode == 0, res.stdout + res.stderr[-3000:]
    for suite, n in (("shutdown_policy_tests", 2), ("pod_names_validation_tests", 1)):
        root = ET.parse(tmp_path / f"junit_{suite}.xml").getroot()
        assert root.attrib["tests"] == str(n) and root.attrib["failures"] == "0"
    assert "ok   ShutdownPolicyTests.test_shutdown_worker0" in res.stdout


def test_all_eight_reference_suites_are_present():
    assert sorted(SUITES) == sorted([
        "simple_tfjob_tests", "shutdown_policy_tests", "cleanpod_policy_tests",
        "replica_restart_policy_tests", "estimator_runconfig_tests", "invalid_tfjob_tests",
        "pod_names_validation_tests", "distributed_training_tests"])
