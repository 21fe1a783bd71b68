"""Job clients: TFJobClient (the reference's SDK surface) and the analogous PyTorchJobClient,
MXJobClient and XGBoostJobClient.

Method set and semantics follow sdk/python/kubeflow/tfjob/api/tf_job_client.py:54-442:
create / get(watch) / patch / delete / wait_for_job / wait_for_condition / get_job_status /
is_job_running / is_job_succeeded / get_pod_names / get_logs; every API error is re-raised as
RuntimeError. The reference has only TFJobClient (sdk/python/kubeflow/tfjob/__init__.py:25);
BASELINE.json asks for a PyTorchJobClient, designed here by analogy (same methods,
``pytorchjobs`` plural).

A client talks to (in this order): ``transport=`` / ``cluster=`` given explicitly, the REST
façade named by ``$TFOP_APISERVER``, or the process-wide default LocalCluster.
"""
from __future__ import annotations

import logging
import os
import queue
import threading
import time
from typing import Any, Callable, Dict, List, Optional

from ..api import module_for
from ..store.memory import ApiError
from . import constants, utils
from .models import Model, to_wire
from .transport import HttpTransport, InProcTransport
from .watch import watch as job_watch

logging.basicConfig(format="%(message)s")
logging.getLogger().setLevel(logging.INFO)


def _wrap_log_stream(q: "queue.Queue", stream) -> None:
    try:
        for line in stream:
            q.put(line)
    finally:
        q.put(None)


def _log_queue_pool(streams) -> List["queue.Queue"]:
    pool = []
    for s in streams:
        q: "queue.Queue" = queue.Queue(maxsize=100)
        pool.append(q)
        threading.Thread(target=_wrap_log_stream, args=(q, s), daemon=True).start()
    return pool


class JobClient:
    KIND = "TFJob"

    def __init__(self, config_file=None, context=None, client_configuration=None,
                 persist_config=True, *, cluster=None, transport=None):
        """Connect to an API server: ``cluster=`` an in-process LocalCluster, ``transport=`` an
        explicit transport, else ``$TFOP_APISERVER`` (REST) or the process-wide default cluster.
        The kubeconfig arguments of the reference signature are accepted and ignored."""
        # config_file / context / client_configuration / persist_config are accepted for
        # signature compatibility with the reference; there is no kubeconfig on one box.
        mod = module_for(self.KIND)
        self.info = mod.INFO
        self.plural = self.info.plural
        if transport is None:
            if cluster is not None:
                transport = InProcTransport(cluster)
            elif os.environ.get("TFOP_APISERVER"):
                transport = HttpTransport(os.environ["TFOP_APISERVER"])
            else:
                from ..cluster import default_cluster
                transport = InProcTransport(default_cluster())
        self.transport = transport

    # ---- helpers -------------------------------------------------------------
    def _body(self, job) -> Dict[str, Any]:
        body = to_wire(job) if isinstance(job, Model) else to_wire(dict(job))
        body.setdefault("apiVersion", f"{constants.TFJOB_GROUP}/{constants.TFJOB_VERSION}")
        body.setdefault("kind", self.KIND)
        return body

    def _err(self, call: str, e: Exception):
        return RuntimeError(f"Exception when calling {call}: {e}\n")

    # ---- CRUD ----------------------------------------------------------------
    def create(self, job, namespace=None):
        """Create the job; returns the created object (a dict)."""
        if namespace is None:
            namespace = utils.set_job_namespace(job)
        try:
            return self.transport.create(self.plural, namespace, self._body(job))
        except ApiError as e:
            raise self._err("CustomObjectsApi->create_namespaced_custom_object", e) from e

    def get(self, name=None, namespace=None, watch=False, timeout_seconds=600):
        """One job (name given), or all jobs of the namespace as ``{"items": [...]}``.
        ``watch=True`` prints a NAME/STATE/TIME table until the job finishes."""
        if namespace is None:
            namespace = utils.get_default_target_namespace()
        try:
            if name:
                if watch:
                    job_watch(self, name=name, namespace=namespace, timeout_seconds=timeout_seconds)
                    return None
                return self.transport.get(self.plural, namespace, name)
            if watch:
                job_watch(self, namespace=namespace, timeout_seconds=timeout_seconds)
                return None
            return {"apiVersion": "kubeflow.org/v1", "kind": self.KIND + "List",
                    "items": self.transport.list(self.plural, namespace)}
        except ApiError as e:
            raise self._err("CustomObjectsApi->get_namespaced_custom_object", e) from e

    def patch(self, name, job, namespace=None):
        """Merge-patch the named job with ``job`` (dict or model); returns the patched object."""
        if namespace is None:
            namespace = utils.set_job_namespace(job)
        try:
            return self.transport.patch(self.plural, namespace, name, self._body(job))
        except ApiError as e:
            raise self._err("CustomObjectsApi->patch_namespaced_custom_object", e) from e

    def delete(self, name, namespace=None):
        """Delete the named job; its pods and services go with it (owner references)."""
        if namespace is None:
            namespace = utils.get_default_target_namespace()
        try:
            return self.transport.delete(self.plural, namespace, name)
        except ApiError as e:
            raise self._err("CustomObjectsApi->delete_namespaced_custom_object", e) from e

    # ---- waiting -------------------------------------------------------------
    def wait_for_job(self, name, namespace=None, timeout_seconds=600, polling_interval=30,
                     watch=False, status_callback: Optional[Callable] = None):
        """Wait for Succeeded or Failed; returns the job."""
        if namespace is None:
            namespace = utils.get_default_target_namespace()
        if watch:
            job_watch(self, name=name, namespace=namespace, timeout_seconds=timeout_seconds)
            return None
        return self.wait_for_condition(name, ["Succeeded", "Failed"], namespace=namespace,
                                       timeout_seconds=timeout_seconds,
                                       polling_interval=polling_interval,
                                       status_callback=status_callback)

    def wait_for_condition(self, name, expected_condition, namespace=None, timeout_seconds=600,
                           polling_interval=30, status_callback: Optional[Callable] = None):
        """Poll ``round(timeout/interval)`` times until ANY condition of the job has one of the
        expected types; RuntimeError on timeout."""
        if namespace is None:
            namespace = utils.get_default_target_namespace()
        job = None
        for _ in range(max(1, round(timeout_seconds / polling_interval))):
            job = self.get(name, namespace=namespace)
            if job:
                if status_callback:
                    status_callback(job)
                conditions = (job.get("status") or {}).get("conditions") or []
                for cond in conditions:
                    if cond.get("type", "") in expected_condition:
                        return job
            time.sleep(polling_interval)
        raise RuntimeError(
            "Timeout waiting for {0} {1} in namespace {2} to enter one of the conditions "
            "{3}.".format(self.KIND, name, namespace, expected_condition), job)

    # ---- status ----------------------------------------------------------------
    def get_job_status(self, name, namespace=None) -> str:
        """Type of the LAST condition ("Running", "Succeeded", ...)."""
        if namespace is None:
            namespace = utils.get_default_target_namespace()
        job = self.get(name, namespace=namespace)
        last = ((job.get("status") or {}).get("conditions") or [{}])[-1]
        return last.get("type", "")

    def is_job_running(self, name, namespace=None) -> bool:
        """True when the last condition is Running."""
        return self.get_job_status(name, namespace=namespace).lower() == "running"

    def is_job_succeeded(self, name, namespace=None) -> bool:
        """True when the last condition is Succeeded."""
        return self.get_job_status(name, namespace=namespace).lower() == "succeeded"

    # ---- pods / logs -------------------------------------------------------------
    def get_pod_names(self, name, namespace=None, master=False, replica_type=None,
                      replica_index=None):
        """Set of pod names of the job, filtered by master role / replica type / replica index
        through the controller's labels; None (with a warning) when nothing matches."""
        if namespace is None:
            namespace = utils.get_default_target_namespace()
        labels = utils.get_labels(name, master=master, replica_type=replica_type,
                                  replica_index=replica_index)
        try:
            pods = self.transport.list("pods", namespace, utils.to_selector(labels))
        except ApiError as e:
            raise self._err("CoreV1Api->list_namespaced_pod", e) from e
        names = [p["metadata"]["name"] for p in pods if p.get("metadata", {}).get("name")]
        if not names:
            logging.warning("Not found Pods of the %s %s with the labels %s.", self.KIND, name, labels)
            return None
        return set(names)

    def get_logs(self, name, namespace=None, master=True, replica_type=None, replica_index=None,
                 follow=False):
        """Log the training output of the job's pods (master-role pod only by default).
        Returns ``{pod: text}`` for the non-follow case as a convenience."""
        if namespace is None:
            namespace = utils.get_default_target_namespace()
        pods = self.get_pod_names(name, namespace=namespace, master=master,
                                  replica_type=replica_type, replica_index=replica_index)
        pod_names = sorted(pods) if pods else []
        if not pod_names:
            raise RuntimeError("Not found Pods of the {} {} in namespace {}".format(
                self.KIND, name, namespace))
        if follow:
            streams = [self.transport.pod_log_stream(namespace, p) for p in pod_names]
            finished = [False] * len(streams)
            pool = _log_queue_pool(streams)
            while not all(finished):
                for i, q in enumerate(pool):
                    if finished[i]:
                        continue
                    for _ in range(50):       # group up to 50 lines of the same pod
                        try:
                            line = q.get(timeout=1)
                        except queue.Empty:
                            break
                        if line is None:
                            finished[i] = True
                            break
                        logging.info("[Pod %s]: %s", pod_names[i], line)
            return None
        out = {}
        for p in pod_names:
            try:
                text = self.transport.pod_log(namespace, p)
            except ApiError as e:
                raise self._err("CoreV1Api->read_namespaced_pod_log", e) from e
            logging.info("The logs of Pod %s:\n %s", p, text)
            out[p] = text
        return out


class TFJobClient(JobClient):
    KIND = "TFJob"


class PyTorchJobClient(JobClient):
    KIND = "PyTorchJob"


class MXJobClient(JobClient):
    KIND = "MXJob"


class XGBoostJobClient(JobClient):
    KIND = "XGBoostJob"
