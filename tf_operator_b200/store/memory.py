"""In-memory Kubernetes-style API server (L0 stand-in for one box).

Implements the subset of kube-apiserver behaviour the operator, launcher and SDK
rely on: namespaced resources, uid / resourceVersion / creationTimestamp
stamping, optimistic concurrency, a /status subresource for the job kinds,
label + field selectors, list/watch with ADDED/MODIFIED/DELETED events,
ownerReference cascade deletion (background and foreground), structural
admission of the four kubeflow.org/v1 CRDs ("spec.tfReplicaSpecs: Required
value", the rejection test/…/invalid_tfjob_tests.py:26-44 expects).

The reference talks to a real apiserver through client-go / controller-runtime
(SURVEY.md layer L0, 2.11); its unit tests fake it with hand-filled informer
caches (pkg/controller.v1/tensorflow/controller_test.go:202-227). This store is
our equivalent of both.
"""
from __future__ import annotations

import collections
import copy
import queue
import threading
import os
import random
import uuid
from typing import Any, Callable, Dict, Iterable, Iterator, List, Optional, Tuple

from ..api import common as c

JOB_RESOURCES = {
    "tfjobs": ("TFJob", "tfReplicaSpecs"),
    "pytorchjobs": ("PyTorchJob", "pytorchReplicaSpecs"),
    "mxjobs": ("MXJob", "mxReplicaSpecs"),
    "xgboostjobs": ("XGBoostJob", "xgbReplicaSpecs"),
}
CORE_KINDS = {
    "pods": "Pod", "services": "Service", "events": "Event", "endpoints": "Endpoints",
    "configmaps": "ConfigMap", "leases": "Lease", "podgroups": "PodGroup",
    "customresourcedefinitions": "CustomResourceDefinition", "namespaces": "Namespace",
}
STATUS_SUBRESOURCE = set(JOB_RESOURCES) | {"pods", "podgroups"}
CLUSTER_SCOPED = {"customresourcedefinitions", "namespaces"}


class ApiError(Exception):
    def __init__(self, code: int, reason: str, message: str):
        super().__init__(message)
        self.code, self.reason, self.message = code, reason, message

    def to_status(self) -> Dict[str, Any]:
        return {"kind": "Status", "apiVersion": "v1", "status": "Failure", "message": self.message,
                "reason": self.reason, "code": self.code}


def not_found(resource, name):
    return ApiError(404, "NotFound", f'{resource} "{name}" not found')


def already_exists(resource, name):
    return ApiError(409, "AlreadyExists", f'{resource} "{name}" already exists')


def conflict(resource, name):
    return ApiError(409, "Conflict",
                    f'Operation cannot be fulfilled on {resource} "{name}": the object has been '
                    f'modified; please apply your changes to the latest version and try again')


def invalid(kind, name, detail):
    return ApiError(422, "Invalid", f'{kind}.kubeflow.org "{name}" is invalid: {detail}')


# ----------------------------------------------------------------------------- selectors
def parse_label_selector(sel) -> List[Tuple[str, str, Any]]:
    if not sel:
        return []
    if isinstance(sel, dict):
        return [(k, "=", v) for k, v in sel.items()]
    out = []
    for part in [p.strip() for p in sel.split(",") if p.strip()]:
        if "!=" in part:
            k, v = part.split("!=", 1)
            out.append((k.strip(), "!=", v.strip()))
        elif "==" in part:
            k, v = part.split("==", 1)
            out.append((k.strip(), "=", v.strip()))
        elif "=" in part:
            k, v = part.split("=", 1)
            out.append((k.strip(), "=", v.strip()))
        elif part.startswith("!"):
            out.append((part[1:].strip(), "!exists", None))
        else:
            out.append((part, "exists", None))
    return out


def labels_match(labels: Optional[Dict[str, str]], reqs) -> bool:
    labels = labels or {}
    for k, op, v in reqs:
        if op == "=" and labels.get(k) != v:
            return False
        if op == "!=" and labels.get(k) == v:
            return False
        if op == "exists" and k not in labels:
            return False
        if op == "!exists" and k in labels:
            return False
    return True


def _field(obj, path: str):
    cur = obj
    for p in path.split("."):
        if not isinstance(cur, dict):
            return None
        cur = cur.get(p)
    return cur


def fields_match(obj, sel: Optional[str]) -> bool:
    if not sel:
        return True
    for part in [p for p in sel.split(",") if p]:
        if "!=" in part:
            k, v = part.split("!=", 1)
            if str(_field(obj, k)) == v:
                return False
        else:
            k, v = part.replace("==", "=").split("=", 1)
            if str(_field(obj, k) or "") != v:
                return False
    return True


# ----------------------------------------------------------------------------- watch
class Watch:
    """Iterator over (type, object) events; ``stop()`` ends it."""

    _STOP = object()

    def __init__(self, store: "ObjectStore", resource: str, namespace: Optional[str],
                 reqs, field_selector):
        self._store, self.resource, self.namespace = store, resource, namespace
        self._reqs, self._fields = reqs, field_selector
        self._q: "queue.Queue" = queue.Queue()
        self._stopped = False

    def _offer(self, etype: str, obj: Dict[str, Any]):
        if self._stopped:
            return
        if self.namespace and c.namespace_of(obj) != self.namespace and \
                self.resource not in CLUSTER_SCOPED:
            return
        if not labels_match(obj.get("metadata", {}).get("labels"), self._reqs):
            return
        if not fields_match(obj, self._fields):
            return
        self._q.put((etype, c.deepcopy(obj)))

    def stop(self):
        self._stopped = True
        self._q.put(self._STOP)
        self._store._remove_watch(self)

    def get(self, timeout: Optional[float] = None):
        try:
            item = self._q.get(timeout=timeout)
        except queue.Empty:
            return None
        return None if item is self._STOP else item

    def __iter__(self) -> Iterator[Tuple[str, Dict[str, Any]]]:
        while True:
            item = self._q.get()
            if item is self._STOP:
                return
            yield item


# ----------------------------------------------------------------------------- store
class ObjectStore:
    def __init__(self):
        self._lock = threading.RLock()
        self._objs: Dict[Tuple[str, str, str], Dict[str, Any]] = {}
        # secondary indexes, maintained wherever _objs changes (_index / _unindex): names by
        # resource, by (resource, namespace, label key, label value) and by owner uid. What a
        # client-go informer's indexers are to the reference's controllers: "the pods of this
        # job" must not cost a walk over every object of the cluster.
        self._by_res: Dict[str, Dict[Tuple[str, str], Dict[str, Any]]] = {}
        self._by_label: Dict[Tuple[str, str, str, str], set] = {}
        self._by_owner: Dict[str, set] = {}
        self._rv = 0
        self._watches: List[Watch] = []
        self._listeners: List[Callable[[str, str, Dict[str, Any]], None]] = []
        # recording hooks, the equivalent of control.FakePodControl (controller_test.go:63-64)
        # (verb, resource, ns, name) of the most recent writes; bounded, as is the number of Event
        # objects kept (kube-apiserver drops events after an hour): a long-lived cluster must not
        # grow with every reconcile it has ever done
        self.audit: "collections.deque[Tuple[str, str, str, str]]" = collections.deque(maxlen=65536)
        self.max_events = 20000
        self._event_order: "collections.deque[Tuple[str, str, str]]" = collections.deque()
        self.admission_enabled = True

    # ---- helpers -----------------------------------------------------------
    def _index(self, key: Tuple[str, str, str], obj: Dict[str, Any]) -> None:
        res, ns, name = key
        self._by_res.setdefault(res, {})[(ns, name)] = obj
        md = obj.get("metadata", {})
        for k, v in (md.get("labels") or {}).items():
            self._by_label.setdefault((res, ns, k, str(v)), set()).add(name)
        for r in md.get("ownerReferences") or []:
            if r.get("uid"):
                self._by_owner.setdefault(r["uid"], set()).add(key)

    def _unindex(self, key: Tuple[str, str, str], obj: Dict[str, Any]) -> None:
        res, ns, name = key
        self._by_res.get(res, {}).pop((ns, name), None)
        md = obj.get("metadata", {})
        for k, v in (md.get("labels") or {}).items():
            names = self._by_label.get((res, ns, k, str(v)))
            if names is not None:
                names.discard(name)
                if not names:
                    del self._by_label[(res, ns, k, str(v))]
        for r in md.get("ownerReferences") or []:
            keys = self._by_owner.get(r.get("uid"))
            if keys is not None:
                keys.discard(key)
                if not keys:
                    del self._by_owner[r["uid"]]

    def select(self, resource: str, namespace: Optional[str], labels: Dict[str, str],
               owner_uid: Optional[str] = None) -> List[Dict[str, Any]]:
        """Copies of the objects of ``resource`` in ``namespace`` that carry ALL of ``labels`` or
        have an owner reference to ``owner_uid`` - through the indexes, i.e. at a cost
        proportional to the answer, not to the store (ControllerRefManager's candidate set)."""
        ns = self._ns(resource, namespace)
        with self._lock:
            names = None
            for k, v in labels.items():
                cur = self._by_label.get((resource, ns, k, str(v)), set())
                names = set(cur) if names is None else names & cur
                if not names:
                    break
            keys = {(resource, ns, n) for n in (names or set())} if labels else set()
            if owner_uid:
                keys |= {k for k in self._by_owner.get(owner_uid, ()) if k[0] == resource and k[1] == ns}
            out = [c.deepcopy(self._objs[k]) for k in keys if k in self._objs]
        out.sort(key=lambda o: (c.namespace_of(o), c.name_of(o)))
        return out

    def _next_rv(self) -> str:
        self._rv += 1
        return str(self._rv)

    @staticmethod
    def _ns(resource: str, namespace: Optional[str]) -> str:
        return "" if resource in CLUSTER_SCOPED else (namespace or "default")

    def _emit(self, resource: str, etype: str, obj: Dict[str, Any]):
        for w in list(self._watches):
            if w.resource == resource:
                w._offer(etype, obj)
        for fn in list(self._listeners):
            try:
                fn(resource, etype, c.deepcopy(obj))
            except Exception:  # a listener must never break the store
                import traceback
                traceback.print_exc()

    def _remove_watch(self, w: Watch):
        with self._lock:
            if w in self._watches:
                self._watches.remove(w)

    def add_listener(self, fn: Callable[[str, str, Dict[str, Any]], None]):
        """fn(resource, event_type, object) is called synchronously for every change."""
        with self._lock:
            self._listeners.append(fn)

    def remove_listener(self, fn):
        with self._lock:
            if fn in self._listeners:
                self._listeners.remove(fn)

    # ---- admission (structural CRD validation + legacy flat spec) -----------
    def _admit_job(self, resource: str, obj: Dict[str, Any]):
        kind, specs_key = JOB_RESOURCES[resource]
        obj.setdefault("apiVersion", c.API_VERSION)
        obj.setdefault("kind", kind)
        if obj["kind"] != kind:
            raise ApiError(400, "BadRequest",
                           f"{obj['kind']} does not match the resource {resource}")
        c.lift_flat_run_policy(obj)
        spec = obj.get("spec")
        name = c.name_of(obj)
        if not isinstance(spec, dict):
            raise invalid(kind, name, "spec: Required value")
        if spec.get(specs_key) is None:
            raise invalid(kind, name, f"spec.{specs_key}: Required value")
        if not isinstance(spec[specs_key], dict):
            raise invalid(kind, name, f"spec.{specs_key}: Invalid value: must be an object")
        if resource == "mxjobs" and not spec.get("jobMode"):
            raise invalid(kind, name, "spec.jobMode: Required value")
        for rt, rs in spec[specs_key].items():
            if rs is None:
                continue
            reps = rs.get("replicas")
            if reps is not None and (not isinstance(reps, int) or isinstance(reps, bool)):
                raise invalid(kind, name, f"spec.{specs_key}.{rt}.replicas: Invalid value: "
                                          f"must be an integer")

    # ---- CRUD ---------------------------------------------------------------
    def create(self, resource: str, obj: Dict[str, Any], namespace: Optional[str] = None) -> Dict[str, Any]:
        obj = c.deepcopy(obj)
        md = obj.setdefault("metadata", {})
        if not md.get("name"):
            if md.get("generateName"):
                md["name"] = md["generateName"] + _new_uid().hex[:5]
            else:
                raise ApiError(422, "Invalid", "metadata.name: Required value")
        ns = self._ns(resource, namespace or md.get("namespace"))
        if ns:
            md["namespace"] = ns
        if resource in JOB_RESOURCES and self.admission_enabled:
            self._admit_job(resource, obj)
        elif resource in CORE_KINDS:
            obj.setdefault("kind", CORE_KINDS[resource])
            obj.setdefault("apiVersion", "v1")
        with self._lock:
            key = (resource, ns, md["name"])
            if key in self._objs:
                raise already_exists(resource, md["name"])
            md["uid"] = md.get("uid") or str(_new_uid())
            md["resourceVersion"] = self._next_rv()
            md["creationTimestamp"] = md.get("creationTimestamp") or c.now_rfc3339()
            md.setdefault("generation", 1)
            if resource == "pods":
                obj.setdefault("status", {}).setdefault("phase", c.POD_PENDING)
            self._objs[key] = obj
            self._index(key, obj)
            if resource == "events":
                self._event_order.append(key)
                while len(self._event_order) > self.max_events:
                    old = self._event_order.popleft()
                    if old in self._objs:
                        self.delete("events", old[1], old[2])
            self.audit.append(("create", resource, ns, md["name"]))
            self._emit(resource, "ADDED", obj)
            return c.deepcopy(obj)

    def get(self, resource: str, namespace: Optional[str], name: str) -> Dict[str, Any]:
        with self._lock:
            obj = self._objs.get((resource, self._ns(resource, namespace), name))
            if obj is None:
                raise not_found(resource, name)
            return c.deepcopy(obj)

    def list(self, resource: str, namespace: Optional[str] = None, label_selector=None,
             field_selector: Optional[str] = None, predicate=None) -> List[Dict[str, Any]]:
        """``predicate(obj) -> bool`` is evaluated on the STORED object (read only!) before the
        copy: a controller that wants "my job's pods" out of thousands copies only those. With
        every reconcile copying every pod of the namespace, 100 jobs took 3.7 s to reach Running
        and 500 did not get there in two minutes (benchmarks/control_plane_scale.py)."""
        reqs = parse_label_selector(label_selector)
        with self._lock:
            out = []
            for (ns, _), obj in self._by_res.get(resource, {}).items():
                if namespace and resource not in CLUSTER_SCOPED and ns != namespace:
                    continue
                if not labels_match(obj.get("metadata", {}).get("labels"), reqs):
                    continue
                if not fields_match(obj, field_selector):
                    continue
                if predicate is not None and not predicate(obj):
                    continue
                out.append(c.deepcopy(obj))
            out.sort(key=lambda o: (c.namespace_of(o), c.name_of(o)))
            return out

    def resource_version(self) -> str:
        with self._lock:
            return str(self._rv)

    def _write(self, resource: str, obj: Dict[str, Any], *, status_only: bool, spec_only: bool):
        md = obj.get("metadata", {})
        ns, name = self._ns(resource, md.get("namespace")), md.get("name")
        with self._lock:
            key = (resource, ns, name)
            cur = self._objs.get(key)
            if cur is None:
                raise not_found(resource, name)
            rv = md.get("resourceVersion")
            if rv and rv != cur["metadata"]["resourceVersion"]:
                raise conflict(resource, name)
            new = c.deepcopy(obj)
            if status_only:
                merged = c.deepcopy(cur)
                merged["status"] = new.get("status", {})
                new = merged
            elif spec_only and resource in STATUS_SUBRESOURCE:
                new["status"] = c.deepcopy(cur.get("status", {}))
            if resource in JOB_RESOURCES and self.admission_enabled and not status_only:
                self._admit_job(resource, new)
            nmd = new.setdefault("metadata", {})
            for k in ("uid", "creationTimestamp", "namespace", "name"):
                if k in cur["metadata"]:
                    nmd[k] = cur["metadata"][k]
            if not status_only and new.get("spec") != cur.get("spec"):
                nmd["generation"] = int(cur["metadata"].get("generation", 1)) + 1
            else:
                nmd["generation"] = cur["metadata"].get("generation", 1)
            if cur["metadata"].get("deletionTimestamp"):
                nmd["deletionTimestamp"] = cur["metadata"]["deletionTimestamp"]
            nmd["resourceVersion"] = self._next_rv()
            self._unindex(key, cur)
            self._objs[key] = new
            self._index(key, new)
            self.audit.append(("update_status" if status_only else "update", resource, ns, name))
            self._emit(resource, "MODIFIED", new)
            return c.deepcopy(new)

    def update(self, resource: str, obj: Dict[str, Any]) -> Dict[str, Any]:
        return self._write(resource, obj, status_only=False, spec_only=True)

    def update_status(self, resource: str, obj: Dict[str, Any]) -> Dict[str, Any]:
        return self._write(resource, obj, status_only=True, spec_only=False)

    def patch(self, resource: str, namespace: Optional[str], name: str,
              patch: Dict[str, Any]) -> Dict[str, Any]:
        """JSON merge patch (RFC 7386)."""
        with self._lock:
            cur = self.get(resource, namespace, name)
            cur["metadata"].pop("resourceVersion", None)
            merged = _merge_patch(cur, patch)
            merged.setdefault("metadata", {})["name"] = name
            return self._write(resource, merged, status_only=False, spec_only=False)

    def delete(self, resource: str, namespace: Optional[str], name: str,
               propagation: str = "Background") -> Dict[str, Any]:
        with self._lock:
            key = (resource, self._ns(resource, namespace), name)
            obj = self._objs.get(key)
            if obj is None:
                raise not_found(resource, name)
            uid = obj["metadata"].get("uid")
            if propagation != "Orphan":
                if propagation == "Foreground":
                    obj["metadata"]["deletionTimestamp"] = c.now_rfc3339()
                # dependents first (foreground) or right after (background): same net effect here
                for (res, ns, nm) in sorted(self._by_owner.get(uid, ())):
                    if (res, ns, nm) in self._objs:
                        self.delete(res, ns, nm, propagation)
            del self._objs[key]
            self._unindex(key, obj)
            obj["metadata"]["resourceVersion"] = self._next_rv()
            obj["metadata"].setdefault("deletionTimestamp", c.now_rfc3339())
            self.audit.append(("delete", resource, key[1], name))
            self._emit(resource, "DELETED", obj)
            return c.deepcopy(obj)

    def watch(self, resource: str, namespace: Optional[str] = None, label_selector=None,
              field_selector: Optional[str] = None, send_initial: bool = True) -> Watch:
        w = Watch(self, resource, namespace, parse_label_selector(label_selector), field_selector)
        with self._lock:
            if send_initial:
                for obj in self.list(resource, namespace):
                    w._offer("ADDED", obj)
            self._watches.append(w)
        return w

    # ---- events (record.EventRecorder) --------------------------------------
    def record_event(self, involved: Dict[str, Any], etype: str, reason: str, message: str,
                     component: str = "training-operator") -> None:
        ns = c.namespace_of(involved)
        ev = {
            "apiVersion": "v1", "kind": "Event",
            "metadata": {"name": f"{c.name_of(involved)}.{_new_uid().hex[:12]}", "namespace": ns},
            "involvedObject": {"kind": involved.get("kind"), "namespace": ns,
                               "name": c.name_of(involved),
                               "uid": involved.get("metadata", {}).get("uid"),
                               "apiVersion": involved.get("apiVersion")},
            "type": etype, "reason": reason, "message": message,
            "source": {"component": component},
            "firstTimestamp": c.now_rfc3339(), "lastTimestamp": c.now_rfc3339(), "count": 1,
        }
        self.create("events", ev, ns)


_uid_rng = random.Random(int.from_bytes(os.urandom(16), "big"))
_uid_lock = threading.Lock()


def _new_uid() -> uuid.UUID:
    """A version-4 UUID from a process-local generator seeded from the OS once. uuid.uuid4() asks
    the kernel for every id; the syscall drops the GIL and, with busy peer threads, getting it back
    costs up to a switch interval - 16 % of the operator's time at 100 concurrent jobs."""
    with _uid_lock:
        return uuid.UUID(int=_uid_rng.getrandbits(128), version=4)


def _merge_patch(target, patch):
    if not isinstance(patch, dict):
        return c.deepcopy(patch)
    if not isinstance(target, dict):
        target = {}
    out = dict(target)
    for k, v in patch.items():
        if v is None:
            out.pop(k, None)
        else:
            out[k] = _merge_patch(out.get(k), v)
    return out
