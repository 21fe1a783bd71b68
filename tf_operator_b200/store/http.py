"""HTTP façade of the in-memory API server: the Kubernetes REST subset that the SDK, the e2e
harness and kubectl-style clients use.

  /apis/kubeflow.org/v1[/namespaces/{ns}]/{tfjobs|pytorchjobs|mxjobs|xgboostjobs}[/{name}[/status]]
  /apis/scheduling.volcano.sh/v1beta1/namespaces/{ns}/podgroups[/{name}]
  /api/v1[/namespaces/{ns}]/{pods|services|events}[/{name}]
  /api/v1/namespaces/{ns}/pods/{name}/log[?container=&follow=true]
  /api/v1/namespaces/{ns}/services/{name}:{port}/proxy/{path}    (what the reference's e2e uses
        to reach test-server: py/kubeflow/tf_operator/util.py:108-139)
  ?watch=true streams {"type","object"} JSON lines; ?labelSelector= / ?fieldSelector= filter.

REST shape encoded by the reference's generated client
(pkg/client/clientset/versioned/typed/tensorflow/v1/tfjob.go:36-48).
"""
from __future__ import annotations

import json
import threading
import urllib.error
import urllib.request
from http.server import BaseHTTPRequestHandler, ThreadingHTTPServer
from typing import Optional
from urllib.parse import parse_qs, urlparse

from ..api import common as c
from .memory import JOB_RESOURCES, ApiError, ObjectStore


def _list_kind(resource: str) -> str:
    if resource in JOB_RESOURCES:
        return JOB_RESOURCES[resource][0] + "List"
    return {"pods": "PodList", "services": "ServiceList", "events": "EventList",
            "podgroups": "PodGroupList"}.get(resource, "List")


class ApiServer:
    def __init__(self, store: ObjectStore, kubelet=None, host: str = "127.0.0.1", port: int = 0):
        self.store, self.kubelet = store, kubelet
        outer = self

        class Handler(BaseHTTPRequestHandler):
            protocol_version = "HTTP/1.1"

            def log_message(self, *a):
                pass

            # ---- plumbing ----
            def _send(self, code: int, obj, ctype="application/json"):
                body = obj if isinstance(obj, bytes) else json.dumps(obj).encode()
                self.send_response(code)
                self.send_header("Content-Type", ctype)
                self.send_header("Content-Length", str(len(body)))
                self.end_headers()
                self.wfile.write(body)

            def _err(self, e: ApiError):
                self._send(e.code, e.to_status())

            def _body(self):
                n = int(self.headers.get("Content-Length") or 0)
                raw = self.rfile.read(n) if n else b""
                return json.loads(raw) if raw else {}

            def _route(self):
                u = urlparse(self.path)
                q = {k: v[0] for k, v in parse_qs(u.query).items()}
                parts = [p for p in u.path.split("/") if p]
                if parts[:1] == ["api"]:
                    parts = parts[2:]
                elif parts[:1] == ["apis"]:
                    parts = parts[3:]
                else:
                    return None, None, None, None, q, parts
                ns = None
                if parts[:1] == ["namespaces"] and len(parts) >= 3:
                    ns, parts = parts[1], parts[2:]
                resource = parts[0] if parts else None
                name = parts[1] if len(parts) > 1 else None
                sub = parts[2:] if len(parts) > 2 else []
                return ns, resource, name, sub, q, parts

            # ---- verbs ----
            def do_GET(self):  # noqa: N802
                try:
                    u = urlparse(self.path)
                    if u.path in ("/healthz", "/readyz", "/livez"):
                        return self._send(200, b"ok", "text/plain")
                    if u.path == "/version":
                        return self._send(200, {"gitVersion": "v0.1.0-b200", "platform": "local"})
                    ns, resource, name, sub, q, _ = self._route()
                    if resource is None:
                        return self._send(404, {"message": "not found"})
                    if resource == "services" and name and ":" in name and sub[:1] == ["proxy"]:
                        return self._proxy(ns, name, "/".join(sub[1:]), u.query)
                    if resource == "pods" and name and sub == ["log"]:
                        return self._logs(ns, name, q)
                    if name is None:
                        if q.get("watch") in ("true", "1"):
                            return self._watch(resource, ns, q)
                        items = outer.store.list(resource, ns, q.get("labelSelector"),
                                                 q.get("fieldSelector"))
                        return self._send(200, {
                            "apiVersion": c.API_VERSION if resource in JOB_RESOURCES else "v1",
                            "kind": _list_kind(resource),
                            "metadata": {"resourceVersion": outer.store.resource_version()},
                            "items": items})
                    return self._send(200, outer.store.get(resource, ns, name))
                except ApiError as e:
                    self._err(e)
                except BrokenPipeError:
                    pass

            def do_POST(self):  # noqa: N802
                try:
                    ns, resource, name, sub, q, _ = self._route()
                    if resource == "services" and name and ":" in name and sub[:1] == ["proxy"]:
                        return self._proxy(ns, name, "/".join(sub[1:]), urlparse(self.path).query)
                    self._send(201, outer.store.create(resource, self._body(), ns))
                except ApiError as e:
                    self._err(e)

            def do_PUT(self):  # noqa: N802
                try:
                    ns, resource, name, sub, q, _ = self._route()
                    obj = self._body()
                    obj.setdefault("metadata", {}).setdefault("namespace", ns)
                    obj["metadata"].setdefault("name", name)
                    if sub == ["status"]:
                        return self._send(200, outer.store.update_status(resource, obj))
                    self._send(200, outer.store.update(resource, obj))
                except ApiError as e:
                    self._err(e)

            def do_PATCH(self):  # noqa: N802
                try:
                    ns, resource, name, sub, q, _ = self._route()
                    self._send(200, outer.store.patch(resource, ns, name, self._body()))
                except ApiError as e:
                    self._err(e)

            def do_DELETE(self):  # noqa: N802
                try:
                    ns, resource, name, sub, q, _ = self._route()
                    body = self._body()
                    prop = body.get("propagationPolicy") or q.get("propagationPolicy") or "Background"
                    outer.store.delete(resource, ns, name, prop)
                    self._send(200, {"kind": "Status", "apiVersion": "v1", "status": "Success"})
                except ApiError as e:
                    self._err(e)

            # ---- streaming ----
            def _chunk(self, data: bytes):
                self.wfile.write(f"{len(data):x}\r\n".encode() + data + b"\r\n")
                self.wfile.flush()

            def _watch(self, resource, ns, q):
                w = outer.store.watch(resource, ns, q.get("labelSelector"), q.get("fieldSelector"))
                self.send_response(200)
                self.send_header("Content-Type", "application/json")
                self.send_header("Transfer-Encoding", "chunked")
                self.end_headers()
                timeout = float(q.get("timeoutSeconds", 0) or 0)
                import time
                end = time.time() + timeout if timeout else None
                try:
                    while True:
                        remaining = None if end is None else end - time.time()
                        if remaining is not None and remaining <= 0:
                            break
                        item = w.get(timeout=min(1.0, remaining) if remaining else 1.0)
                        if item is None:
                            continue
                        et, obj = item
                        self._chunk(json.dumps({"type": et, "object": obj}).encode() + b"\n")
                except (BrokenPipeError, ConnectionResetError):
                    pass
                finally:
                    w.stop()
                    try:
                        self._chunk(b"")
                    except OSError:
                        pass

            def _logs(self, ns, name, q):
                if outer.kubelet is None:
                    return self._send(501, {"message": "no kubelet attached"})
                outer.store.get("pods", ns, name)   # 404 if the pod is gone
                container = q.get("container")
                if q.get("follow") in ("true", "1"):
                    self.send_response(200)
                    self.send_header("Content-Type", "text/plain")
                    self.send_header("Transfer-Encoding", "chunked")
                    self.end_headers()
                    try:
                        for line in outer.kubelet.follow_log(ns, name, container):
                            self._chunk((line + "\n").encode())
                        self._chunk(b"")
                    except (BrokenPipeError, ConnectionResetError):
                        pass
                    return
                self._send(200, outer.kubelet.read_log(ns, name, container).encode(), "text/plain")

            def _proxy(self, ns, svc_port, path, query):
                from ..launcher import netmap
                svc, _, port = svc_port.partition(":")
                ip = netmap.pod_ip(ns, svc, getattr(outer.kubelet.opt, "cluster_id", None))
                url = f"http://{ip}:{port}/{path}" + (f"?{query}" if query else "")
                try:
                    with urllib.request.urlopen(url, timeout=10) as r:
                        self._send(r.status, r.read(), r.headers.get("Content-Type", "text/plain"))
                except urllib.error.HTTPError as e:
                    self._send(e.code, e.read() or b"", "text/plain")
                except Exception as e:  # noqa: BLE001 - the target exited while answering /exit
                    self._send(503, {"message": f"proxy error: {e}"})

        self._srv = ThreadingHTTPServer((host, port), Handler)
        self._srv.daemon_threads = True
        self.host, self.port = self._srv.server_address[:2]
        self._thread: Optional[threading.Thread] = None

    @property
    def url(self) -> str:
        return f"http://{self.host}:{self.port}"

    def start(self) -> "ApiServer":
        self._thread = threading.Thread(target=self._srv.serve_forever, daemon=True,
                                        name="tfop-apiserver")
        self._thread.start()
        return self

    def stop(self) -> None:
        self._srv.shutdown()
        self._srv.server_close()
