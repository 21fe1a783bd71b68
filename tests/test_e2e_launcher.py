"""End-to-end control-plane suites against the single-box cluster (API server + operator +
launcher), driven through the SDK and the fault-injecting test-server.

Python re-expression of the reference's Argo/EKS e2e suites:
simple_tfjob_tests.py, shutdown_policy_tests.py, cleanpod_policy_tests.py,
replica_restart_policy_tests.py, estimator_runconfig_tests.py, pod_names_validation_tests.py,
invalid_tfjob_tests.py — replicas are the same kind of fake workload (examples/test_server).
"""
import json
import os
import sys
import time
import urllib.request

import pytest
import yaml

from tf_operator_b200.api import common as c
from tf_operator_b200.cluster import LocalCluster
from tf_operator_b200.launcher import netmap
from tf_operator_b200.launcher.kubelet import KubeletOptions
from tf_operator_b200.launcher.topology import synthetic
from tf_operator_b200.sdk import PyTorchJobClient, TFJobClient

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TEST_APP = ["python", os.path.join(REPO, "examples", "test_server", "test_app.py")]


@pytest.fixture()
def cluster():
    cl = LocalCluster(topology=synthetic(0), serve_http=True,
                      kubelet_options=KubeletOptions(cwd=REPO, restart_backoff=0.05))
    yield cl
    cl.stop()


def tfjob(name, replicas, policy="Never", clean=None, extra=None):
    specs = {rt: {"replicas": n, "restartPolicy": policy, "template": {"spec": {"containers": [
        {"name": "tensorflow", "image": "test-server:local", "command": TEST_APP}]}}}
        for rt, n in replicas.items()}
    job = {"apiVersion": "kubeflow.org/v1", "kind": "TFJob", "metadata": {"name": name, "namespace": "default"},
           "spec": {"tfReplicaSpecs": specs}}
    if clean:
        job["spec"]["runPolicy"] = {"cleanPodPolicy": clean}
    if extra:
        job["spec"].update(extra)
    return job


def wait(pred, timeout=30.0, interval=0.02):
    end = time.time() + timeout
    while time.time() < end:
        v = pred()
        if v:
            return v
        time.sleep(interval)
    raise AssertionError("timed out waiting for condition")


def http_get(cluster, replica, path, port=2222, retries=200):
    """Through the API server's service proxy, like the reference harness (util.py:108-139)."""
    url = f"{cluster.api.url}/api/v1/namespaces/default/services/{replica}:{port}/proxy/{path}"
    last = None
    for _ in range(retries):
        try:
            with urllib.request.urlopen(url, timeout=5) as r:
                return r.read().decode()
        except Exception as e:  # noqa: BLE001 - server not listening yet
            last = e
            time.sleep(0.05)
    raise AssertionError(f"GET {url} failed: {last}")


def wait_running(cluster, name, n):
    def ok():
        pods = cluster.store.list("pods", "default", f"job-name={name}")
        return len(pods) == n and all(p["status"]["phase"] == "Running" for p in pods)
    wait(ok)
    for p in cluster.store.list("pods", "default", f"job-name={name}"):
        http_get(cluster, c.name_of(p), "")     # the replica's server is accepting connections


def phases(cluster, name):
    return {c.name_of(p): p["status"]["phase"] for p in cluster.store.list("pods", "default", f"job-name={name}")}


# simple_tfjob_tests.py:26-90 — run, succeed, delete, re-create under the same name --------------
def test_simple_job_two_trials(cluster):
    cli = TFJobClient(cluster=cluster)
    for trial in range(2):
        cli.create(tfjob("simple", {"Chief": 1, "PS": 1, "Worker": 2}))
        wait_running(cluster, "simple", 4)
        assert cli.is_job_running("simple")
        http_get(cluster, "simple-chief-0", "exit?exitCode=0")
        cli.wait_for_job("simple", polling_interval=0.05, timeout_seconds=30)
        assert cli.is_job_succeeded("simple")
        cli.delete("simple")
        wait(lambda: not cluster.store.list("pods", "default", "job-name=simple"))
        assert cluster.store.list("services", "default", "job-name=simple") == []
    evs = [e["message"] for e in cluster.store.list("events", "default")]
    assert sum("Created pod: simple-" in m for m in evs) == 8
    assert sum("Created service: simple-" in m for m in evs) == 8


# pod_names_validation_tests.py:14,47-84 ----------------------------------------------------------
def test_pod_and_service_names(cluster):
    cli = TFJobClient(cluster=cluster)
    cli.create(tfjob("names", {"Chief": 1, "PS": 2, "Worker": 3}))
    wait_running(cluster, "names", 6)
    expected = {"names-chief-0", "names-ps-0", "names-ps-1", "names-worker-0", "names-worker-1", "names-worker-2"}
    assert cli.get_pod_names("names") == expected
    assert {c.name_of(s) for s in cluster.store.list("services", "default")} == expected
    assert cli.get_pod_names("names", master=True) == {"names-chief-0"}
    assert cli.get_pod_names("names", replica_type="ps", replica_index=1) == {"names-ps-1"}


# shutdown_policy_tests.py:25-93 ------------------------------------------------------------------
@pytest.mark.parametrize("replicas,kill", [({"Chief": 1, "PS": 1, "Worker": 2}, "chief-0"),
                                           ({"PS": 1, "Worker": 2}, "worker-0")])
def test_shutdown_policy(cluster, replicas, kill):
    cli = TFJobClient(cluster=cluster)
    cli.create(tfjob("shut", replicas))
    wait_running(cluster, "shut", sum(replicas.values()))
    http_get(cluster, f"shut-{kill}", "exit?exitCode=0")
    cli.wait_for_job("shut", polling_interval=0.05, timeout_seconds=30)
    assert cli.is_job_succeeded("shut")


# cleanpod_policy_tests.py:26-115 ----------------------------------------------------------------
@pytest.mark.parametrize("policy,remaining", [
    ("All", set()), ("Running", {"clean-chief-0"}), ("None", {"clean-chief-0", "clean-ps-0", "clean-worker-0"})])
def test_clean_pod_policy(cluster, policy, remaining):
    cli = TFJobClient(cluster=cluster)
    cli.create(tfjob("clean", {"Chief": 1, "PS": 1, "Worker": 1}, clean=policy))
    wait_running(cluster, "clean", 3)
    http_get(cluster, "clean-chief-0", "exit?exitCode=0")
    cli.wait_for_job("clean", polling_interval=0.05, timeout_seconds=30)
    wait(lambda: set(phases(cluster, "clean")) == remaining)
    time.sleep(0.2)
    assert set(phases(cluster, "clean")) == remaining


# replica_restart_policy_tests.py:27-156 -----------------------------------------------------------
def _start_time(cluster, pod):
    st = cluster.store.get("pods", "default", pod)["status"]["containerStatuses"][0]
    return st["restartCount"], st["state"]


@pytest.mark.parametrize("policy,code,restarts,final", [
    ("Always", 0, True, None), ("Always", 1, True, None),
    ("OnFailure", 1, True, None), ("OnFailure", 0, False, "Succeeded"),
    ("Never", 1, False, "Failed"), ("Never", 0, False, "Succeeded"),
    ("ExitCode", 1, False, "Failed"), ("ExitCode", 130, True, None),
])
def test_replica_restart_policy(cluster, policy, code, restarts, final):
    cli = TFJobClient(cluster=cluster)
    cli.create(tfjob("rp", {"Worker": 1, "PS": 1}, policy=policy))
    wait_running(cluster, "rp", 2)
    uid0 = cluster.store.get("pods", "default", "rp-worker-0")["metadata"]["uid"]
    assert cluster.store.get("pods", "default", "rp-worker-0")["spec"]["restartPolicy"] == \
        ("Never" if policy == "ExitCode" else policy)
    http_get(cluster, "rp-worker-0", f"exit?exitCode={code}")
    if restarts:
        if policy == "ExitCode":
            # the operator deletes the failed pod and re-creates the same name/index
            wait(lambda: cluster.store.list("pods", "default", "job-name=rp,replica-type=worker") and
                 cluster.store.get("pods", "default", "rp-worker-0")["metadata"]["uid"] != uid0)
            job = cli.get("rp")
            assert not c.is_failed(job["status"])
        else:
            # kubelet-style in-place restart: same pod object, restartCount grows
            wait(lambda: cluster.store.get("pods", "default", "rp-worker-0")["status"]
                 ["containerStatuses"][0]["restartCount"] >= 1)
            assert cluster.store.get("pods", "default", "rp-worker-0")["metadata"]["uid"] == uid0
        wait_running(cluster, "rp", 2)
        assert cli.get_job_status("rp") in ("Running", "Restarting")
    else:
        cli.wait_for_job("rp", polling_interval=0.05, timeout_seconds=30)
        assert cli.get_job_status("rp") == final


# estimator_runconfig_tests.py:25-101 ----------------------------------------------------------------
def test_estimator_runconfig(cluster):
    cli = TFJobClient(cluster=cluster)
    cli.create(tfjob("rc", {"Chief": 1, "PS": 2, "Worker": 2, "Evaluator": 1}))
    wait_running(cluster, "rc", 6)
    tfc = json.loads(http_get(cluster, "rc-worker-1", "tfconfig"))
    assert tfc["task"] == {"type": "worker", "index": 1} and tfc["environment"] == "cloud"
    assert tfc["cluster"]["ps"] == ["rc-ps-0.default.svc:2222", "rc-ps-1.default.svc:2222"]
    for replica, ttype, tid, chief in (("rc-chief-0", "chief", 0, True), ("rc-worker-0", "worker", 0, False),
                                       ("rc-worker-1", "worker", 1, False), ("rc-ps-1", "ps", 1, False)):
        rc = json.loads(http_get(cluster, replica, "runconfig"))
        assert rc["task_type"] == ttype and rc["task_id"] == tid and rc["is_chief"] is chief
        assert rc["master"] == f"grpc://{replica}.default.svc:2222"
        assert rc["num_worker_replicas"] == 3 and rc["num_ps_replicas"] == 2
        assert set(rc["cluster_spec"]) == {"chief", "ps", "worker"}
    ev = json.loads(http_get(cluster, "rc-evaluator-0", "runconfig"))
    assert ev["cluster_spec"] == {} and ev["task_type"] == "evaluator" and ev["master"] == ""
    assert ev["num_worker_replicas"] == 0 and ev["num_ps_replicas"] == 0


# invalid_tfjob_tests.py:26-44 -------------------------------------------------------------------------
def test_invalid_tfjob_is_rejected_by_the_api_server(cluster):
    cli = TFJobClient(cluster=cluster)
    with pytest.raises(RuntimeError) as e:
        cli.create({"apiVersion": "kubeflow.org/v1", "kind": "TFJob", "metadata": {"name": "invalid"},
                    "spec": {"notTheActualField": {}}})
    assert "spec.tfReplicaSpecs: Required value" in str(e.value)


def test_legacy_invalid_spec_marks_job_failed():
    """Legacy tf-operator.v1 behaviour (job.go:62-112): a spec that fails validation gets a
    Failed condition with reason InvalidTFJobSpec written to /status."""
    from tf_operator_b200.operator.manager import ManagerOptions
    cl = LocalCluster(topology=synthetic(0), manager_options=ManagerOptions(
        enabled_schemes=["TFJob"], fail_invalid_spec=True))
    try:
        bad = tfjob("bad", {"Worker": 1})
        bad["spec"]["tfReplicaSpecs"]["Worker"]["template"]["spec"]["containers"][0]["name"] = "main"
        cl.store.create("tfjobs", bad)
        def failed():
            st = cl.store.get("tfjobs", "default", "bad").get("status") or {}
            cd = c.get_condition(st, "Failed") if st.get("conditions") else None
            return cd
        cd = wait(failed)
        assert cd["reason"] == "InvalidTFJobSpec"
        assert cl.store.list("pods") == []
    finally:
        cl.stop()


# gang admission (pod.go:220-237 + Volcano semantics) -----------------------------------------------
def test_gang_admission_is_all_or_nothing():
    from tf_operator_b200.operator.manager import ManagerOptions
    cl = LocalCluster(topology=synthetic(2), serve_http=True,
                      manager_options=ManagerOptions(enable_gang_scheduling=True),
                      kubelet_options=KubeletOptions(cwd=REPO, gpus_per_replica_default=1))
    try:
        cli = TFJobClient(cluster=cl)
        cli.create(tfjob("big", {"Worker": 3}))         # wants 3 GPUs, box has 2
        time.sleep(0.5)
        assert set(phases(cl, "big").values()) == {"Pending"}
        assert cl.kubelet.free_gpus() == [0, 1]          # nothing was claimed partially
        assert cl.store.get("podgroups", "default", "big")["spec"]["minMember"] == 3
        cli.create(tfjob("fits", {"Worker": 2}))
        wait_running(cl, "fits", 2)
        assert cl.kubelet.free_gpus() == []
        envs = [json.loads(http_get(cl, f"fits-worker-{i}", "env")) for i in range(2)]
        assert sorted(e["CUDA_VISIBLE_DEVICES"] for e in envs) == ["0", "1"]
        assert set(phases(cl, "big").values()) == {"Pending"}
    finally:
        cl.stop()


def test_pod_that_does_not_fit_waits_for_a_gpu_and_then_runs():
    """A pod that finds no free GPU stays Pending (PodScheduled=False, Unschedulable) and is placed
    as soon as another pod gives its GPU back - what kube-scheduler's retry queue does. The case
    behind it: a job re-created right after its same-named predecessor was deleted, while one of
    the old replicas was still exiting."""
    cl = LocalCluster(topology=synthetic(1), serve_http=True,
                      kubelet_options=KubeletOptions(cwd=REPO, gpus_per_replica_default=1))
    try:
        cli = TFJobClient(cluster=cl)
        cli.create(tfjob("first", {"Worker": 1}))
        wait_running(cl, "first", 1)
        cli.create(tfjob("second", {"Worker": 1}))
        pod = wait(lambda: [p for p in cl.store.list("pods", "default", "job-name=second")
                            if (p.get("status") or {}).get("conditions")])[0]
        cond = pod["status"]["conditions"][0]
        assert (pod["status"]["phase"], cond["type"], cond["status"], cond["reason"]) == \
            ("Pending", "PodScheduled", "False", "Unschedulable")
        assert cl.kubelet.free_gpus() == []
        http_get(cl, "first-worker-0", "exit?exitCode=0")
        wait_running(cl, "second", 1)
        env = json.loads(http_get(cl, "second-worker-0", "env"))
        assert env["CUDA_VISIBLE_DEVICES"] == "0"
        http_get(cl, "second-worker-0", "exit?exitCode=0")
        cli.wait_for_job("second", polling_interval=0.05, timeout_seconds=30)
        assert cli.is_job_succeeded("second") and cl.kubelet.free_gpus() == [0]
    finally:
        cl.stop()


# BASELINE.json config 1 ---------------------------------------------------------------------------------
def test_dist_mnist_tfjob_1ps_2workers_cpu(cluster):
    job = yaml.safe_load(open(os.path.join(REPO, "examples", "dist_mnist", "tf_job_mnist.yaml")))
    for spec in job["spec"]["tfReplicaSpecs"].values():
        cmd = spec["template"]["spec"]["containers"][0]["command"]
        spec["template"]["spec"]["containers"][0]["command"] = [
            a.replace("--train_steps=200", "--train_steps=60") for a in cmd]
    cli = TFJobClient(cluster=cluster)
    cli.create(job)
    done = cli.wait_for_job("dist-mnist-for-e2e-test", polling_interval=0.2, timeout_seconds=120)
    assert cli.is_job_succeeded("dist-mnist-for-e2e-test"), done["status"]
    logs = cli.get_logs("dist-mnist-for-e2e-test", master=True)
    text = logs["dist-mnist-for-e2e-test-worker-0"]
    assert "validation cross entropy" in text and "job name = worker" in text
    # PS never exits by itself; cleanPodPolicy=Running (the TFJob default) removed it
    wait(lambda: "dist-mnist-for-e2e-test-ps-0" not in phases(cluster, "dist-mnist-for-e2e-test"))


def test_pytorchjob_env_rendezvous_gloo(cluster):
    """PyTorchJob: MASTER_ADDR/RANK/WORLD_SIZE contract drives a real gloo all-reduce."""
    script = ("import torch, torch.distributed as dist;"
              "from tf_operator_b200.runtime.env import ReplicaContext;"
              "ctx = ReplicaContext.from_env(); ctx.init_process_group('gloo');"
              "t = torch.ones(4) * (ctx.rank + 1); dist.all_reduce(t);"
              "print('rank', ctx.rank, 'world', ctx.world_size, 'sum', t[0].item())")
    spec = lambda n: {"replicas": n, "restartPolicy": "Never", "template": {"spec": {"containers": [  # noqa: E731
        {"name": "pytorch", "image": "x", "command": ["python", "-c", script]}]}}}
    job = {"apiVersion": "kubeflow.org/v1", "kind": "PyTorchJob", "metadata": {"name": "ptg", "namespace": "default"},
           "spec": {"pytorchReplicaSpecs": {"Master": spec(1), "Worker": spec(2)}}}
    cli = PyTorchJobClient(cluster=cluster)
    cli.create(job)
    cli.wait_for_job("ptg", polling_interval=0.2, timeout_seconds=120)
    assert cli.is_job_succeeded("ptg")
    logs = cli.get_logs("ptg", master=False)
    assert "rank 0 world 3 sum 6.0" in logs["ptg-master-0"]
    assert "rank 2 world 3 sum 6.0" in logs["ptg-worker-1"]


# the reference's other example workloads, run as jobs ------------------------
def _run_manifest(cluster, path, client_cls, overrides=None, timeout=150):
    job = yaml.safe_load(open(os.path.join(REPO, path)))
    if overrides:
        overrides(job)
    cli = client_cls(cluster=cluster)
    cli.create(job)
    name = job["metadata"]["name"]
    cli.wait_for_job(name, polling_interval=0.2, timeout_seconds=timeout)
    return cli, name


def test_tf_smoke_job_chief_reaches_every_task(cluster):
    cli, name = _run_manifest(cluster, "examples/tf_sample/tf_job_smoke.yaml", TFJobClient)
    assert cli.is_job_succeeded(name)
    log = cli.get_logs(name, master=True)[f"{name}-chief-0"]
    for task in ("/job:chief/task:0", "/job:ps/task:1", "/job:worker/task:3"):
        assert task + " ok" in log
    assert "all tasks answered" in log


def test_estimator_style_collective_allreduce_job(cluster):
    cli, name = _run_manifest(cluster, "examples/distribution_strategy/estimator_api/distributed_tfjob.yaml",
                              TFJobClient)
    assert cli.is_job_succeeded(name)
    logs = cli.get_logs(name, master=False)
    assert len(logs) == 3 and all("eval loss" in t for t in logs.values())


def test_keras_style_multi_worker_cnn_checkpoints_and_resumes(cluster, tmp_path):
    def cpu(job):
        ctr = job["spec"]["tfReplicaSpecs"]["Worker"]["template"]["spec"]["containers"][0]
        ctr.pop("resources")
        ctr["command"] = ctr["command"][:2] + ["--epochs", "1", "--steps-per-epoch", "30",
                                               "--checkpoint_dir", str(tmp_path / "ckpt"),
                                               "--saved_model_dir", str(tmp_path / "saved")]
        ctr["env"] = [{"name": "CUDA_VISIBLE_DEVICES", "value": ""}]
    cli, name = _run_manifest(cluster, "examples/distribution_strategy/keras_api/multi_worker_tfjob.yaml",
                              TFJobClient, cpu)
    assert cli.is_job_succeeded(name), cli.get_logs(name, master=False)
    assert (tmp_path / "ckpt" / "ckpt.pt").exists() and (tmp_path / "saved" / "model.pt").exists()
    cli.delete(name)
    wait(lambda: not cluster.store.list("pods", "default", f"job-name={name}"))

    def resume(job):
        cpu(job)
        ctr = job["spec"]["tfReplicaSpecs"]["Worker"]["template"]["spec"]["containers"][0]
        ctr["command"][ctr["command"].index("--epochs") + 1] = "2"
    cli, name = _run_manifest(cluster, "examples/distribution_strategy/keras_api/multi_worker_tfjob.yaml",
                              TFJobClient, resume)
    assert cli.is_job_succeeded(name)
    assert "resumed from" in cli.get_logs(name, master=True)[f"{name}-worker-0"]


def test_mnist_with_summaries_job(cluster, tmp_path):
    def small(job):
        ctr = job["spec"]["tfReplicaSpecs"]["Worker"]["template"]["spec"]["containers"][0]
        ctr["command"] = ctr["command"][:2] + [f"--log_dir={tmp_path}", "--max_steps=120", "--learning_rate=0.01"]
    cli, name = _run_manifest(cluster, "examples/mnist_with_summaries/tf_job_mnist.yaml", TFJobClient, small)
    assert cli.is_job_succeeded(name)
    events = [json.loads(l) for l in open(tmp_path / "test" / "events.jsonl")]
    assert events[-1]["accuracy"] > 0.8 and "layer1/weights" in events[0]
    assert any("trace_ms" in json.loads(l) for l in open(tmp_path / "train" / "events.jsonl"))


def test_mxjob_kvstore_training(cluster):
    from tf_operator_b200.sdk import MXJobClient
    cli, name = _run_manifest(cluster, "examples/mxnet/mxjob_dist_v1.yaml", MXJobClient)
    assert cli.is_job_succeeded(name)
    log = cli.get_logs(name, master=False, replica_type="worker")[f"{name}-worker-0"]
    assert "DMLC_ROLE=worker servers=1 workers=1 root=mxnet-job-scheduler-0:9091" in log and "final loss" in log
    wait(lambda: set(phases(cluster, name)) == {f"{name}-worker-0"})     # cleanPodPolicy Running


# replica_restart_policy_tests.py:27-156 with the data-plane life cycle of the GPU job, on CPU -------
@pytest.mark.parametrize("world", [2, 8])
def test_killed_replica_reforms_the_gang_control_plane_version(world, tmp_path):
    """The kill-and-restart cycle of tests/test_multigpu_gpu.py with the GPU parts faked
    (tests/workers/fake_ddp.py): rank 1 is SIGKILLed at step 25 (exit 137, retryable), its peers
    lose it, exit 138, the operator re-creates every replica, the gang re-forms in a later
    rendezvous generation - whatever order the restarts arrive in - and resumes from step 20."""
    cl = LocalCluster(topology=synthetic(world), serve_http=False,
                      kubelet_options=KubeletOptions(cwd=REPO, restart_backoff=0.05,
                                                     gpus_per_replica_default=1))
    try:
        ckpt, marker, res = tmp_path / "ckpt.json", tmp_path / "killed", tmp_path / "result.json"
        cmd = ["python", os.path.join(REPO, "tests", "workers", "fake_ddp.py"), "40", str(ckpt), str(res), "2.0"]

        def spec(n):
            ctr = {"name": "pytorch", "image": "test:local", "command": cmd,
                   "env": [{"name": "TFOP_FAULT_INJECT", "value": f"1:25:{marker}"},
                           {"name": "FAKE_START_JITTER_S", "value": os.environ.get("FAKE_START_JITTER_S", "0.5")}]}
            return {"replicas": n, "restartPolicy": "ExitCode", "template": {"spec": {"containers": [ctr]}}}
        job = {"apiVersion": "kubeflow.org/v1", "kind": "PyTorchJob",
               "metadata": {"name": "fake-restart", "namespace": "default"},
               "spec": {"runPolicy": {"cleanPodPolicy": "None"},
                        "pytorchReplicaSpecs": {"Master": spec(1), "Worker": spec(world - 1)}}}
        cli = PyTorchJobClient(cluster=cl)
        cli.create(job)
        cli.wait_for_job("fake-restart", polling_interval=0.05, timeout_seconds=300)
        logs = cli.get_logs("fake-restart", master=False)
        dump = "\n".join(f"--- {k}\n{v[-1500:]}" for k, v in logs.items())
        assert cli.is_job_succeeded("fake-restart"), dump
        assert marker.exists()
        out = json.load(open(res))
        assert out["world"] == world and out["generation"] >= 1 and out["resumed_from"] == 20, out
    finally:
        cl.stop()
