"""Test-side TFJob helpers (reference: py/kubeflow/tf_operator/tf_job_client.py:23-470 - the
e2e harness's own client, distinct from the SDK). Functions take the LocalCluster where the
reference takes a kubernetes ApiClient; names, defaults' meaning and log lines are kept.
"""
from __future__ import annotations

import datetime
import json
import logging
import time
from typing import Callable, Iterable, List, Optional

from ..api import common as c
from ..store.memory import ApiError
from . import k8s_util, util

TF_JOB_GROUP = "kubeflow.org"
TF_JOB_PLURAL = "tfjobs"
TF_JOB_KIND = "TFJob"
TF_JOB_NAME_LABEL = "job-name"        # tf_job_client.py:20-29
JOB_NAME_LABEL = "job-name"

# the reference polls every 30 s for up to 30 min against a real cluster; one box is quicker
DEFAULT_POLL = 0.05


def _try_get(cluster, namespace: str, name: str) -> Optional[dict]:
    try:
        return cluster.store.get(TF_JOB_PLURAL, namespace, name)
    except ApiError as e:
        if e.code == 404:
            return None
        raise


def create_tf_job(cluster, spec: dict, version: str = "v1") -> dict:
    """tf_job_client.py:32-70."""
    ns = spec.get("metadata", {}).get("namespace", "default")
    api_response = cluster.store.create(TF_JOB_PLURAL, spec)
    logging.info("Created job %s in namespace %s", c.name_of(api_response), ns)
    return api_response


def delete_tf_job(cluster, namespace: str, name: str, version: str = "v1") -> dict:
    """Foreground deletion: dependants go with the job (tf_job_client.py:73-103)."""
    logging.info("Deleting job %s.%s", namespace, name)
    return cluster.store.delete(TF_JOB_PLURAL, namespace, name, propagation="Foreground")


def log_status(tf_job: dict) -> None:
    """tf_job_client.py:117-126."""
    conds = tf_job.get("status", {}).get("conditions") or []
    logging.info("Job %s in namespace %s; uid=%s; conditions=%s", tf_job["metadata"].get("name"),
                 tf_job["metadata"].get("namespace"), tf_job["metadata"].get("uid"),
                 [(x.get("type"), x.get("status")) for x in conds])


def wait_for_condition(cluster, namespace: str, name: str, expected_condition: Iterable[str],
                       version: str = "v1",
                       timeout: datetime.timedelta = datetime.timedelta(minutes=2),
                       polling_interval: float = DEFAULT_POLL,
                       status_callback: Optional[Callable[[dict], None]] = None) -> dict:
    """Wait until any condition of a type in expected_condition is True
    (tf_job_client.py:129-196)."""
    expected = set(expected_condition)
    end = time.time() + timeout.total_seconds()
    while True:
        job = _try_get(cluster, namespace, name)
        if job:
            if status_callback:
                status_callback(job)
            for cond in job.get("status", {}).get("conditions") or []:
                if cond.get("type") in expected and cond.get("status") == "True":
                    return job
        if time.time() > end:
            raise util.TimeoutError_(
                f"Timeout waiting for job {name} in namespace {namespace} to enter one of the "
                f"conditions {sorted(expected)}.")
        time.sleep(polling_interval)


def wait_for_job(cluster, namespace: str, name: str, version: str = "v1",
                 timeout: datetime.timedelta = datetime.timedelta(minutes=2),
                 polling_interval: float = DEFAULT_POLL,
                 status_callback: Optional[Callable[[dict], None]] = None) -> dict:
    """Wait for Succeeded or Failed (tf_job_client.py:199-222)."""
    return wait_for_condition(cluster, namespace, name, ["Succeeded", "Failed"], version, timeout,
                              polling_interval, status_callback)


def wait_for_delete(cluster, namespace: str, name: str, version: str = "v1",
                    timeout: datetime.timedelta = datetime.timedelta(minutes=2),
                    polling_interval: float = DEFAULT_POLL,
                    status_callback: Optional[Callable[[dict], None]] = None) -> None:
    """tf_job_client.py:225-266."""
    end = time.time() + timeout.total_seconds()
    while True:
        job = _try_get(cluster, namespace, name)
        if job is None:
            return
        if status_callback:
            status_callback(job)
        if time.time() > end:
            raise util.TimeoutError_(f"Timeout waiting for job {name} in namespace {namespace} to be deleted.")
        time.sleep(polling_interval)


def get_labels(name: str, replica_type: Optional[str] = None, replica_index: Optional[int] = None) -> dict:
    """Labels the controller puts on a job's pods / services (tf_job_client.py:269-283)."""
    labels = {"group-name": TF_JOB_GROUP, JOB_NAME_LABEL: name}
    if replica_type:
        labels["replica-type"] = str(replica_type).lower()
    if replica_index is not None:
        labels["replica-index"] = str(replica_index)
    return labels


def to_selector(labels: dict) -> str:
    return ",".join(f"{k}={v}" for k, v in labels.items())


def get_pod_names(cluster, namespace: str, name: str) -> set:
    """tf_job_client.py:293-311."""
    pods = k8s_util.list_pods(cluster.store, namespace, to_selector(get_labels(name)))
    return {c.name_of(p) for p in pods}


def wait_for_replica_type_in_phases(cluster, namespace: str, tfjob_name: str, replica_type: str,
                                    phases: Iterable[str],
                                    timeout: datetime.timedelta = datetime.timedelta(minutes=2)) -> List[dict]:
    """tf_job_client.py:314-329."""
    selector = to_selector(get_labels(tfjob_name, replica_type))
    return k8s_util.wait_for_pods_to_be_in_phases(cluster.store, namespace, selector, phases, timeout)


def terminate_replica(cluster, namespace: str, target: str, exit_code: int = 0, port: int = 2222) -> None:
    """Ask one replica's test-server to exit with exit_code, through the API server's service
    proxy (tf_job_client.py:358-377)."""
    util.send_request(cluster.api.url, namespace, target, "exit", {"exitCode": exit_code}, port)


def terminate_replicas(cluster, namespace: str, name: str, replica: str, num_targets: int,
                       exit_code: int = 0) -> None:
    """Terminate replicas 0..num_targets-1 of a type once they run (tf_job_client.py:332-356)."""
    wait_for_replica_type_in_phases(cluster, namespace, name, replica, ["Running"])
    for i in range(num_targets):
        terminate_replica(cluster, namespace, f"{name}-{replica.lower()}-{i}", exit_code)


def job_succeeded(tf_job: dict) -> bool:
    """Last condition is Succeeded (tf_job_client.py:380-397)."""
    conds = tf_job.get("status", {}).get("conditions") or []
    return bool(conds) and conds[-1].get("type", "").lower() == "succeeded"


def get_creation_failures_from_tfjob(cluster, namespace: str, tfjob: dict) -> List[str]:
    """Compare the 'Created pod/service' events with the replica counts in the spec
    (tf_job_client.py:400-444): returns human-readable discrepancies, [] when all were created."""
    uid = tfjob["metadata"].get("uid")
    events = k8s_util.get_events(cluster.store, uid, namespace)
    created_pods, created_services = k8s_util.parse_events(events)
    num_expected = sum(int(spec.get("replicas", 1))
                       for spec in tfjob["spec"].get("tfReplicaSpecs", {}).values())
    failures = []
    if len(created_pods) != num_expected:
        failures.append(f"Expected {num_expected} pods to be created but only got {len(created_pods)} "
                        "create events.")
    if len(created_services) != num_expected:
        failures.append(f"Expected {num_expected} services to be created but only got "
                        f"{len(created_services)} create events.")
    return failures


def get_start_time_by_index(cluster, namespace: str, name: str, replica_type: str, replica_index: int,
                            phase: str) -> Optional[str]:
    """tf_job_client.py:447-455."""
    selector = to_selector(get_labels(name, replica_type))
    return k8s_util.get_container_start_time(cluster.store, namespace, selector, replica_index, [phase])


def terminate_and_verify_start_time(cluster, namespace: str, name: str, replica: str, index: int,
                                    exit_code: int, expect_restart: bool) -> bool:
    """Kill one replica and report whether its container was started again as the restart policy
    says it should (tf_job_client.py:458-470). The reference compares container start times; on
    one box a restart can land within the same RFC3339 second, so the pod uid (ExitCode policy:
    the operator re-creates the pod) and restartCount (kubelet restarts in place) count too."""
    selector = to_selector(get_labels(name, replica, index))

    def snapshot():
        pods = [p for p in k8s_util.list_pods(cluster.store, namespace, selector)
                if p.get("status", {}).get("phase") == "Running"]
        if not pods:
            return None
        p = pods[0]
        cs = (p.get("status", {}).get("containerStatuses") or [{}])[0]
        started = (cs.get("state", {}).get("running") or {}).get("startedAt")
        return p["metadata"].get("uid"), started, cs.get("restartCount", 0)

    first = None
    end = time.time() + 30
    while first is None and time.time() < end:
        first = snapshot()
        time.sleep(0.02)
    terminate_replica(cluster, namespace, f"{name}-{replica.lower()}-{index}", exit_code)
    if not expect_restart:
        time.sleep(0.5)
        now = snapshot()
        return now is None or now == first
    end = time.time() + 30
    while time.time() < end:
        now = snapshot()
        if now is not None and now != first:
            return True
        time.sleep(0.05)
    return False


def dump(tf_job: dict) -> str:
    return json.dumps(tf_job, indent=1, default=str)
