"""Cluster-side helpers of the e2e harness (reference: py/kubeflow/tf_operator/k8s_util.py:14-223).

Every function takes the object store (``cluster.store``) where the reference takes a
kubernetes ``ApiClient``; selectors, phases and the event-message regex are the reference's.
"""
from __future__ import annotations

import datetime
import logging
import re
import time
from typing import Dict, Iterable, List, Optional

from ..api import common as c

# k8s_util.py:140 - events of the form "Created pod: name" / "Created service: name"
_CREATED_RE = re.compile(r".*Created.*(pod|[Ss]ervice).*: (.*)")


def list_pods(store, namespace: str, label_selector: str) -> List[dict]:
    """k8s_util.py:104-121."""
    return store.list("pods", namespace, label_selector)


def wait_for_pods_to_be_in_phases(store, namespace: str, pod_selector: str, phases: Iterable[str],
                                  timeout: datetime.timedelta = datetime.timedelta(minutes=2),
                                  polling_interval: float = 0.05) -> List[dict]:
    """Block until at least one pod matches and every matching pod is in one of `phases`
    (k8s_util.py:14-60)."""
    phases = set(phases)
    end = time.time() + timeout.total_seconds()
    while True:
        pods = list_pods(store, namespace, pod_selector)
        if pods and all(p.get("status", {}).get("phase") in phases for p in pods):
            return pods
        if time.time() > end:
            seen = {c.name_of(p): p.get("status", {}).get("phase") for p in pods}
            raise TimeoutError(f"timeout waiting for pods {pod_selector!r} to be in {sorted(phases)}: {seen}")
        time.sleep(polling_interval)


def wait_for_pods_to_be_deleted(store, namespace: str, pod_selector: str,
                                timeout: datetime.timedelta = datetime.timedelta(minutes=2),
                                polling_interval: float = 0.05) -> None:
    """k8s_util.py:63-101."""
    end = time.time() + timeout.total_seconds()
    while True:
        pods = list_pods(store, namespace, pod_selector)
        if not pods:
            logging.info("no pods matching %s found", pod_selector)
            return
        if time.time() > end:
            raise TimeoutError(f"timeout waiting for pods {pod_selector!r} to be deleted")
        time.sleep(polling_interval)


def get_events(store, uid: str, namespace: Optional[str] = None) -> List[dict]:
    """Events whose involvedObject has this uid (k8s_util.py:124-152)."""
    return [e for e in store.list("events", namespace)
            if e.get("involvedObject", {}).get("uid") == uid]


def parse_events(events: Iterable[dict]):
    """-> (created pod names, created service names) from the event messages
    (k8s_util.py:155-184)."""
    pods, services = set(), set()
    for e in events:
        m = _CREATED_RE.match(e.get("message", ""))
        if not m:
            continue
        (pods if m.group(1) == "pod" else services).add(m.group(2).strip())
    return pods, services


def get_container_start_time(store, namespace: str, pod_selector: str, index: int,
                             phases: Iterable[str]) -> Optional[str]:
    """startedAt of the first container of the index-th matching pod, once it is in `phases`
    (k8s_util.py:187-223)."""
    pods = wait_for_pods_to_be_in_phases(store, namespace, pod_selector, phases)
    pods = sorted(pods, key=c.name_of)
    if index >= len(pods):
        return None
    for cs in pods[index].get("status", {}).get("containerStatuses", []):
        state = cs.get("state", {})
        for key in ("running", "terminated"):
            if key in state and state[key].get("startedAt"):
                return state[key]["startedAt"]
    return None


def replica_phase_map(store, namespace: str, job_name: str) -> Dict[str, str]:
    return {c.name_of(p): p.get("status", {}).get("phase")
            for p in list_pods(store, namespace, f"job-name={job_name}")}
