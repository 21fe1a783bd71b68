"""Pod / Service / PodGroup control: the write side of the reconcile engine.

``Real*`` write to an ObjectStore and emit the create/delete Events the e2e harness
parses (py/kubeflow/tf_operator/k8s_util.py:206: ``.*Created.*(pod|Service).*: (.*)``).
``Fake*`` record what would have been written, like ``control.FakePodControl`` /
``FakeServiceControl`` in the reference's unit tests
(pkg/controller.v1/tensorflow/controller_test.go:63-64, pod_test.go:204).
"""
from __future__ import annotations

import copy
import threading
from typing import Any, Dict, List, Optional

from ..api import common as c
from ..store.memory import ApiError, ObjectStore

SUCCESSFUL_CREATE_POD = "SuccessfulCreatePod"
FAILED_CREATE_POD = "FailedCreatePod"
SUCCESSFUL_DELETE_POD = "SuccessfulDeletePod"
FAILED_DELETE_POD = "FailedDeletePod"
SUCCESSFUL_CREATE_SERVICE = "SuccessfulCreateService"
FAILED_CREATE_SERVICE = "FailedCreateService"
SUCCESSFUL_DELETE_SERVICE = "SuccessfulDeleteService"
FAILED_DELETE_SERVICE = "FailedDeleteService"


class Recorder:
    """record.EventRecorder over the store."""

    def __init__(self, store: Optional[ObjectStore], component: str):
        self.store, self.component = store, component
        self.events: List[str] = []          # "<type> <reason> <message>", like FakeRecorder
        self._lock = threading.Lock()

    def event(self, obj, etype: str, reason: str, message: str):
        with self._lock:
            self.events.append(f"{etype} {reason} {message}")
            if len(self.events) > 10000:
                del self.events[:5000]
        if self.store is not None:
            try:
                self.store.record_event(obj, etype, reason, message, self.component)
            except ApiError:
                pass

    def eventf(self, obj, etype, reason, fmt, *args):
        self.event(obj, etype, reason, fmt % args if args else fmt)


def pod_from_template(template: Dict[str, Any], namespace: str, job, controller_ref) -> Dict[str, Any]:
    tmd = template.get("metadata") or {}
    pod = {
        "apiVersion": "v1", "kind": "Pod",
        "metadata": {"name": tmd.get("name"), "namespace": namespace,
                     "labels": c.deepcopy(tmd.get("labels") or {}),
                     "annotations": c.deepcopy(tmd.get("annotations") or {}),
                     "ownerReferences": [c.deepcopy(controller_ref)] if controller_ref else []},
        "spec": c.deepcopy(template.get("spec") or {}),
        "status": {"phase": c.POD_PENDING},
    }
    if not pod["metadata"]["annotations"]:
        del pod["metadata"]["annotations"]
    if tmd.get("generateName") and not tmd.get("name"):
        pod["metadata"]["generateName"] = tmd["generateName"]
    return pod


class RealPodControl:
    def __init__(self, store: ObjectStore, recorder: Recorder):
        self.store, self.recorder = store, recorder

    def create_pods_with_controller_ref(self, namespace, template, job, controller_ref):
        pod = pod_from_template(template, namespace, job, controller_ref)
        if not pod["metadata"].get("labels"):
            raise ValueError("unable to create pods, no labels")
        try:
            new = self.store.create("pods", pod, namespace)
        except ApiError as e:
            self.recorder.eventf(job, "Warning", FAILED_CREATE_POD, "Error creating: %s", e.message)
            raise
        self.recorder.eventf(job, "Normal", SUCCESSFUL_CREATE_POD, "Created pod: %s", c.name_of(new))
        return new

    def delete_pod(self, namespace, name, job):
        try:
            self.store.delete("pods", namespace, name)
        except ApiError as e:
            if e.code == 404:
                return
            self.recorder.eventf(job, "Warning", FAILED_DELETE_POD, "Error deleting: %s", e.message)
            raise
        self.recorder.eventf(job, "Normal", SUCCESSFUL_DELETE_POD, "Deleted pod: %s", name)

    def patch_pod(self, namespace, name, patch):
        return self.store.patch("pods", namespace, name, patch)


class RealServiceControl:
    def __init__(self, store: ObjectStore, recorder: Recorder):
        self.store, self.recorder = store, recorder

    def create_services_with_controller_ref(self, namespace, service, job, controller_ref):
        svc = c.deepcopy(service)
        svc.setdefault("apiVersion", "v1")
        svc.setdefault("kind", "Service")
        svc.setdefault("metadata", {})["ownerReferences"] = [c.deepcopy(controller_ref)]
        try:
            new = self.store.create("services", svc, namespace)
        except ApiError as e:
            self.recorder.eventf(job, "Warning", FAILED_CREATE_SERVICE, "Error creating: %s", e.message)
            raise
        self.recorder.eventf(job, "Normal", SUCCESSFUL_CREATE_SERVICE, "Created service: %s",
                             c.name_of(new))
        return new

    def delete_service(self, namespace, name, job):
        try:
            self.store.delete("services", namespace, name)
        except ApiError as e:
            if e.code == 404:
                return
            self.recorder.eventf(job, "Warning", FAILED_DELETE_SERVICE, "Error deleting: %s", e.message)
            raise
        self.recorder.eventf(job, "Normal", SUCCESSFUL_DELETE_SERVICE, "Deleted service: %s", name)

    def patch_service(self, namespace, name, patch):
        return self.store.patch("services", namespace, name, patch)


class FakePodControl:
    """Records Templates / ControllerRefs / DeletePodName; ``err`` injects a failure."""

    def __init__(self):
        self.templates: List[Dict[str, Any]] = []
        self.controller_refs: List[Dict[str, Any]] = []
        self.delete_pod_name: List[str] = []
        self.patches: List[Any] = []
        self.err: Optional[Exception] = None
        self.create_limit = 0
        self.create_call_count = 0
        self._lock = threading.Lock()

    def create_pods_with_controller_ref(self, namespace, template, job, controller_ref):
        with self._lock:
            self.create_call_count += 1
            if self.create_limit and self.create_call_count > self.create_limit:
                raise RuntimeError("not creating pod, limit %d already reached" % self.create_limit)
            if self.err is not None:
                raise self.err
            self.templates.append(c.deepcopy(template))
            self.controller_refs.append(c.deepcopy(controller_ref))
        return pod_from_template(template, namespace, job, controller_ref)

    def delete_pod(self, namespace, name, job):
        with self._lock:
            if self.err is not None:
                raise self.err
            self.delete_pod_name.append(name)

    def patch_pod(self, namespace, name, patch):
        with self._lock:
            self.patches.append((namespace, name, patch))

    def clear(self):
        self.__init__()


class FakeServiceControl:
    def __init__(self):
        self.services: List[Dict[str, Any]] = []
        self.controller_refs: List[Dict[str, Any]] = []
        self.delete_service_name: List[str] = []
        self.patches: List[Any] = []
        self.err: Optional[Exception] = None
        self._lock = threading.Lock()

    def create_services_with_controller_ref(self, namespace, service, job, controller_ref):
        with self._lock:
            if self.err is not None:
                raise self.err
            self.services.append(c.deepcopy(service))
            self.controller_refs.append(c.deepcopy(controller_ref))
        return service

    def delete_service(self, namespace, name, job):
        with self._lock:
            if self.err is not None:
                raise self.err
            self.delete_service_name.append(name)

    def patch_service(self, namespace, name, patch):
        with self._lock:
            self.patches.append((namespace, name, patch))

    def clear(self):
        self.__init__()
