"""Pluggable transports under the SDK clients: in-process (a LocalCluster) or HTTP (the REST
façade, store/http.py). Both expose the handful of calls the reference SDK makes through the
``kubernetes`` package (CustomObjectsApi + CoreV1Api; sdk/python/kubeflow/tfjob/api/
tf_job_client.py:72-74), which is not installed here and is not needed.
"""
from __future__ import annotations

import json
import urllib.error
import urllib.parse
import urllib.request
from typing import Any, Dict, Iterator, List, Optional, Tuple

from ..store.memory import ApiError


class InProcTransport:
    def __init__(self, cluster):
        self.cluster = cluster
        self.store = cluster.store

    def create(self, plural, namespace, body): return self.store.create(plural, body, namespace)
    def get(self, plural, namespace, name): return self.store.get(plural, namespace, name)
    def list(self, plural, namespace, label_selector=None):
        return self.store.list(plural, namespace, label_selector)
    def patch(self, plural, namespace, name, body): return self.store.patch(plural, namespace, name, body)
    def delete(self, plural, namespace, name, propagation="Background"):
        return self.store.delete(plural, namespace, name, propagation)

    def watch(self, plural, namespace, timeout_seconds: float) -> Iterator[Tuple[str, Dict[str, Any]]]:
        import time
        w = self.store.watch(plural, namespace)
        end = time.time() + timeout_seconds
        try:
            while time.time() < end:
                item = w.get(timeout=min(0.5, max(0.0, end - time.time())))
                if item is not None:
                    yield item
        finally:
            w.stop()

    def pod_log(self, namespace, pod, container=None) -> str:
        self.store.get("pods", namespace, pod)
        return self.cluster.kubelet.read_log(namespace, pod, container)

    def pod_log_stream(self, namespace, pod, container=None) -> Iterator[str]:
        return self.cluster.kubelet.follow_log(namespace, pod, container)


class HttpTransport:
    def __init__(self, base_url: str, timeout: float = 120.0):
        self.base = base_url.rstrip("/")
        self.timeout = timeout

    @staticmethod
    def _path(plural, namespace, name=None, sub=None):
        core = plural in ("pods", "services", "events")
        root = "/api/v1" if core else (
            "/apis/scheduling.volcano.sh/v1beta1" if plural == "podgroups" else "/apis/kubeflow.org/v1")
        p = root + (f"/namespaces/{namespace}" if namespace else "") + f"/{plural}"
        if name:
            p += f"/{name}"
        if sub:
            p += f"/{sub}"
        return p

    def _req(self, method, path, body=None, query=None, raw=False):
        url = self.base + path + ("?" + urllib.parse.urlencode(query) if query else "")
        data = json.dumps(body).encode() if body is not None else None
        req = urllib.request.Request(url, data=data, method=method,
                                     headers={"Content-Type": "application/json"})
        try:
            with urllib.request.urlopen(req, timeout=self.timeout) as r:
                payload = r.read()
        except urllib.error.HTTPError as e:
            try:
                st = json.loads(e.read())
            except Exception:  # noqa: BLE001
                st = {}
            raise ApiError(e.code, st.get("reason", "Unknown"), st.get("message", str(e))) from None
        return payload.decode() if raw else (json.loads(payload) if payload else {})

    def create(self, plural, namespace, body): return self._req("POST", self._path(plural, namespace), body)
    def get(self, plural, namespace, name): return self._req("GET", self._path(plural, namespace, name))
    def list(self, plural, namespace, label_selector=None):
        q = {"labelSelector": label_selector} if label_selector else None
        return self._req("GET", self._path(plural, namespace), query=q).get("items", [])
    def patch(self, plural, namespace, name, body):
        return self._req("PATCH", self._path(plural, namespace, name), body)
    def delete(self, plural, namespace, name, propagation="Background"):
        return self._req("DELETE", self._path(plural, namespace, name),
                         {"propagationPolicy": propagation})

    def watch(self, plural, namespace, timeout_seconds: float):
        url = self.base + self._path(plural, namespace) + \
            f"?watch=true&timeoutSeconds={int(timeout_seconds)}"
        with urllib.request.urlopen(url, timeout=timeout_seconds + 5) as r:
            for line in r:
                line = line.strip()
                if line:
                    ev = json.loads(line)
                    yield ev["type"], ev["object"]

    def pod_log(self, namespace, pod, container=None) -> str:
        q = {"container": container} if container else None
        return self._req("GET", self._path("pods", namespace, pod, "log"), query=q, raw=True)

    def pod_log_stream(self, namespace, pod, container=None) -> Iterator[str]:
        url = self.base + self._path("pods", namespace, pod, "log") + "?follow=true" + \
            (f"&container={container}" if container else "")
        with urllib.request.urlopen(url, timeout=self.timeout) as r:
            for line in r:
                yield line.decode(errors="replace").rstrip("\n")
