"""Job fixtures of the e2e harness: the 15 ksonnet job components of
test/workflows/components/*.jsonnet as plain builders ``COMPONENTS[name](name, namespace)``.

The replica image of the reference (tf-operator-test-server) becomes the same fake workload run
as a process: ``examples/test_server/test_app.py`` (/tfconfig, /runconfig, /exit?exitCode=N).
``ubuntu`` containers running ``echo Hello`` / ``tail -f /dev/null`` keep their commands.
"""
from __future__ import annotations

import os
import sys
from typing import Callable, Dict, Optional

REPO = os.path.dirname(os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
TEST_SERVER_CMD = [sys.executable, os.path.join(REPO, "examples", "test_server", "test_app.py")]
_ANNOT = {"sidecar.istio.io/inject": "false"}


def _replica(replicas: int, policy: str, command=None, gpu: int = 0, image: str = "test-server:local"):
    ctr = {"name": "tensorflow", "image": image, "command": list(command or TEST_SERVER_CMD)}
    if gpu:
        ctr["resources"] = {"limits": {"nvidia.com/gpu": gpu}}
    return {"replicas": replicas, "restartPolicy": policy,
            "template": {"metadata": {"annotations": dict(_ANNOT)}, "spec": {"containers": [ctr]}}}


def _job(name: str, namespace: str, specs: dict, clean: Optional[str] = None) -> dict:
    spec: dict = {"tfReplicaSpecs": specs}
    if clean is not None:
        spec["cleanPodPolicy"] = clean      # the v1 fixtures use the flat (pre-RunPolicy) field
    return {"apiVersion": "kubeflow.org/v1", "kind": "TFJob",
            "metadata": {"name": name, "namespace": namespace}, "spec": spec}


def simple_tfjob_v1(name, namespace="default"):
    return _job(name, namespace, {"Chief": _replica(1, "Never"), "PS": _replica(2, "Never"),
                                  "Worker": _replica(4, "Never")})


def gpu_tfjob_v1(name, namespace="default"):
    """One Chief asking for one GPU, OnFailure workers (gpu_tfjob_v1.jsonnet:18-60)."""
    return _job(name, namespace, {"Chief": _replica(1, "Never", gpu=1),
                                  "Worker": _replica(1, "OnFailure", gpu=1)})


def master_is_chief_v1(name, namespace="default"):
    return _job(name, namespace, {"Chief": _replica(1, "Never"), "PS": _replica(2, "Never"),
                                  "Worker": _replica(4, "Never")})


def worker0_is_chief_v1(name, namespace="default"):
    return _job(name, namespace, {"PS": _replica(2, "Never"), "Worker": _replica(4, "Never")})


def _clean_pod(policy):
    def build(name, namespace="default"):
        echo, tail = ["echo", "Hello"], ["tail", "-f", "/dev/null"]
        return _job(name, namespace, {"Chief": _replica(1, "Never", echo, image="ubuntu"),
                                      "PS": _replica(2, "Never", tail, image="ubuntu"),
                                      "Worker": _replica(4, "Never", echo, image="ubuntu")}, clean=policy)
    build.__name__ = f"clean_pod_{policy.lower()}_v1"
    return build


clean_pod_all_v1 = _clean_pod("All")
clean_pod_running_v1 = _clean_pod("Running")
clean_pod_none_v1 = _clean_pod("None")


def _restart_policy(policy):
    def build(name, namespace="default"):
        return _job(name, namespace, {"PS": _replica(1, policy), "Worker": _replica(2, policy)})
    build.__name__ = f"replica_restart_policy_{policy.lower()}_v1"
    return build


replica_restart_policy_always_v1 = _restart_policy("Always")
replica_restart_policy_onfailure_v1 = _restart_policy("OnFailure")
replica_restart_policy_never_v1 = _restart_policy("Never")
replica_restart_policy_exitcode_v1 = _restart_policy("ExitCode")


def estimator_runconfig_v1(name, namespace="default"):
    return _job(name, namespace, {"Chief": _replica(1, "Never"), "PS": _replica(2, "Never"),
                                  "Worker": _replica(2, "Never"), "Evaluator": _replica(1, "Never")},
                clean="All")


def distributed_training_v1(name, namespace="default"):
    """3 workers running a real collective training loop (distributed_training_v1.jsonnet:18-36;
    the reference image runs a MirroredStrategy/Keras script - ours is the repo's example)."""
    cmd = [sys.executable, os.path.join(REPO, "examples", "distribution_strategy", "estimator_api",
                                        "collective_allreduce.py")]
    return _job(name, namespace, {"Worker": _replica(3, "Never", cmd, image="distributed-training:local")})


def invalid_tfjob_v1(name, namespace="default"):
    """spec.notTheActualField instead of tfReplicaSpecs (invalid_tfjob_v1.jsonnet:31-47)."""
    tmpl = {"spec": {"containers": [{"name": "tensorflow", "image": "busybox"}]}}
    return {"apiVersion": "kubeflow.org/v1", "kind": "TFJob",
            "metadata": {"name": name, "namespace": namespace},
            "spec": {"notTheActualField": {
                "Ps": {"replicas": 2, "restartPolicy": "Never", "template": tmpl},
                "Worker": {"replicas": 4, "restartPolicy": "Never", "template": tmpl}}}}


def pod_names_validation_v1(name, namespace="default"):
    return _job(name, namespace, {"Chief": _replica(1, "Never"), "PS": _replica(2, "Never"),
                                  "Worker": _replica(3, "Never"), "Evaluator": _replica(1, "Never")},
                clean="All")


COMPONENTS: Dict[str, Callable[..., dict]] = {f.__name__: f for f in [
    simple_tfjob_v1, gpu_tfjob_v1, master_is_chief_v1, worker0_is_chief_v1, clean_pod_all_v1,
    clean_pod_running_v1, clean_pod_none_v1, replica_restart_policy_always_v1,
    replica_restart_policy_onfailure_v1, replica_restart_policy_never_v1,
    replica_restart_policy_exitcode_v1, estimator_runconfig_v1, distributed_training_v1,
    invalid_tfjob_v1, pod_names_validation_v1]}


def build(component: str, name: Optional[str] = None, namespace: str = "default") -> dict:
    """`ks param set <component> name <name>` + `ks show` of the reference, in one call."""
    return COMPONENTS[component](name or component.replace("_", "-"), namespace)
