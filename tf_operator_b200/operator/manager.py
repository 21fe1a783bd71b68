"""The operator process: watches the four job kinds and their pods/services, queues job keys,
and runs the per-kind reconcilers on worker threads.

Covers both reference binaries:
  * training-operator (cmd/training-operator.v1/main.go:58-124): all kinds, --enable-scheme,
    metrics :8080, probes :8081, --leader-elect.
  * tf-operator.v1 (cmd/tf-operator.v1/app/server.go:72-195, controller.go:193-286):
    rate-limited workqueue + `threadiness` workers, namespace scoping, gang scheduling flag,
    leader election (lease 15 s / renew 5 s / retry 3 s), CRD existence check, invalid specs
    marked Failed.
Unlike the reference's new manager (FakeWorkQueue: AddAfter is a no-op, SURVEY.md Appendix
B.1) the deadline / TTL timers here actually fire.
"""
from __future__ import annotations

import logging
import threading
import time
import uuid
from dataclasses import dataclass, field
from http.server import BaseHTTPRequestHandler, ThreadingHTTPServer
from typing import Dict, Iterable, List, Optional

from ..api import common as c
from ..engine.job_controller import JobControllerConfig
from ..engine.workqueue import RateLimitingQueue
from ..kinds import RECONCILERS
from ..store.memory import ApiError, ObjectStore
from . import pprof
from .metrics import OperatorMetrics

log = logging.getLogger("tf_operator_b200.operator")

LEASE_DURATION, RENEW_DEADLINE, RETRY_PERIOD = 15.0, 5.0, 3.0   # server.go:54-59



class EnabledSchemes(list):
    """Repeatable, case-insensitive --enable-scheme (register_controller.go:51-76)."""

    def set(self, kind: str) -> None:
        for k in RECONCILERS:
            if k.lower() == kind.lower():
                if k not in self:
                    self.append(k)
                return
        raise ValueError(f"scheme {kind} is not supported, supported: {sorted(RECONCILERS)}")

    def fill_all(self) -> None:
        for k in RECONCILERS:
            if k not in self:
                self.append(k)

    def empty(self) -> bool:
        return len(self) == 0


@dataclass
class ManagerOptions:
    enabled_schemes: List[str] = field(default_factory=list)
    namespace: str = ""                  # "" = all namespaces
    threadiness: int = 1
    enable_gang_scheduling: bool = False
    gang_scheduler_name: str = c.GANG_SCHEDULER_NAME
    leader_elect: bool = False
    leader_election_id: str = "1ca428e5.training-operator"   # main.go:84
    identity: str = ""
    resync_period: float = 15.0          # JobController reconcile period (controller.go:128)
    metrics_bind_address: str = ""       # ":8080" in the reference; "" disables
    health_probe_bind_address: str = ""  # ":8081"
    fail_invalid_spec: bool = False      # legacy tf-operator.v1 behaviour
    check_crd: bool = False              # legacy: refuse to start without the TFJob CRD
    qps: float = 0.0                     # >0: client-go style write throttle (legacy default 5)
    burst: int = 10


class Manager:
    def __init__(self, store: ObjectStore, options: Optional[ManagerOptions] = None,
                 metrics: Optional[OperatorMetrics] = None):
        self.store = store
        self.opt = options or ManagerOptions()
        schemes = EnabledSchemes()
        for s in self.opt.enabled_schemes:
            schemes.set(s)
        if schemes.empty():
            schemes.fill_all()
        self.schemes = list(schemes)
        self.metrics = metrics or OperatorMetrics()
        self.identity = self.opt.identity or f"operator-{uuid.uuid4().hex[:8]}"
        cfg = JobControllerConfig(reconciler_sync_loop_period=self.opt.resync_period,
                                  enable_gang_scheduling=self.opt.enable_gang_scheduling,
                                  gang_scheduler_name=self.opt.gang_scheduler_name)
        self.queues: Dict[str, RateLimitingQueue] = {}
        self.reconcilers = {}
        for kind in self.schemes:
            q = RateLimitingQueue(name=kind)
            r = RECONCILERS[kind](store, config=cfg, workqueue=q, metrics=self.metrics)
            r.fail_invalid_spec = self.opt.fail_invalid_spec
            if self.opt.qps > 0:
                from ..engine.ratelimit import ThrottledControl, TokenBucket
                if not hasattr(self, "_bucket"):
                    self._bucket = TokenBucket(self.opt.qps, self.opt.burst)
                r.pod_control = ThrottledControl(r.pod_control, self._bucket)
                r.service_control = ThrottledControl(r.service_control, self._bucket)
                _orig = r.update_job_status_in_api_server

                def _throttled_status(job, status, _orig=_orig):
                    self._bucket.accept()
                    return _orig(job, status)
                r.update_job_status_in_api_server = _throttled_status
            self.queues[kind] = q
            self.reconcilers[kind] = r
        self._by_plural = {r.info.plural: k for k, r in self.reconcilers.items()}
        self._threads: List[threading.Thread] = []
        self._stop = threading.Event()
        self._leader = threading.Event()
        self._servers: List[ThreadingHTTPServer] = []
        self._last_seen: Dict[tuple, dict] = {}
        self._created_at: Dict[str, float] = {}
        self._last_ads: Dict[str, object] = {}
        self.started = False

    # ------------------------------------------------------------------ events
    def _in_scope(self, obj) -> bool:
        return not self.opt.namespace or c.namespace_of(obj) == self.opt.namespace

    def _on_event(self, resource: str, etype: str, obj) -> None:
        if not self._in_scope(obj):
            return
        kind = self._by_plural.get(resource)
        if kind is not None:
            key = c.key_of(obj)
            if etype == "ADDED":
                self._created_at.setdefault(f"{kind}:{key}", time.time())
                self._last_ads[f"{kind}:{key}"] = c.run_policy(obj).get("activeDeadlineSeconds")
            if etype == "DELETED":
                self._created_at.pop(f"{kind}:{key}", None)
                self._last_ads.pop(f"{kind}:{key}", None)
                for rt in c.replica_specs(obj, self.reconcilers[kind].info):
                    from ..engine import expectations as exp
                    e = self.reconcilers[kind].expectations
                    e.delete_expectations(exp.gen_expectation_pods_key(key, rt))
                    e.delete_expectations(exp.gen_expectation_services_key(key, rt))
                return
            if etype == "MODIFIED":
                # re-arm the deadline timer when activeDeadlineSeconds is edited on a started job
                # (updateTFJob, pkg/controller.v1/tensorflow/job.go:160-175)
                ads = c.run_policy(obj).get("activeDeadlineSeconds")
                start = (obj.get("status") or {}).get("startTime")
                prev = self._last_ads.get(f"{kind}:{key}")
                if ads is not None and start and prev is not None and prev != ads:
                    remaining = float(ads) - (time.time() - c.parse_time(start))
                    self.queues[kind].add_after(key, max(0.0, remaining))
                self._last_ads[f"{kind}:{key}"] = ads
            if etype == "MODIFIED" and c.has_condition(c.status_of(obj), c.JOB_RUNNING):
                t0 = self._created_at.pop(f"{kind}:{key}", None)
                if t0 is not None:
                    self.metrics.launch_to_running_seconds.labels(kind).observe(time.time() - t0)
            self.queues[kind].add(key)
            return
        if resource in ("pods", "services"):
            ref = c.get_controller_of(obj)
            if ref is None or ref.get("kind") not in self.reconcilers:
                return
            r = self.reconcilers[ref["kind"]]
            ident = (resource, c.namespace_of(obj), c.name_of(obj))
            if etype == "ADDED":
                job_key = r.on_dependent_create(resource, obj)
                self._last_seen[ident] = obj
            elif etype == "DELETED":
                job_key = r.on_dependent_delete(resource, obj)
                self._last_seen.pop(ident, None)
            else:
                job_key = r.on_dependent_update(resource, self._last_seen.get(ident), obj)
                self._last_seen[ident] = obj
            if job_key:
                self.queues[ref["kind"]].add(job_key)

    # ------------------------------------------------------------------ workers
    def _worker(self, kind: str) -> None:
        q, r = self.queues[kind], self.reconcilers[kind]
        while not self._stop.is_set():
            if self.opt.leader_elect and not self._leader.wait(0.2):
                continue
            key = q.get(timeout=0.2)
            if key is None:
                continue
            t0 = time.perf_counter()
            try:
                ns, name = key.split("/", 1)
                r.reconcile(ns, name)
                q.forget(key)
            except Exception:  # noqa: BLE001 - a failed sync is retried with back-off
                log.exception("error syncing %s %s", kind, key)
                q.add_rate_limited(key)
            finally:
                q.done(key)
                dt = time.perf_counter() - t0
                self.metrics.reconcile_seconds.labels(kind).observe(dt)
                # controller.go:303-307: "Finished syncing tfjob %q (%v)"
                log.info("Finished syncing %s %r (%.3fms)", kind.lower(), key, dt * 1e3)

    def _resync(self) -> None:
        while not self._stop.wait(self.opt.resync_period):
            for kind, r in self.reconcilers.items():
                for job in self.store.list(r.info.plural, self.opt.namespace or None):
                    self.queues[kind].add(c.key_of(job))

    # ------------------------------------------------------------------ leader election
    def _elect(self) -> None:
        ns = "kubeflow"
        name = self.opt.leader_election_id
        while not self._stop.is_set():
            now = time.time()
            try:
                lease = self.store.get("leases", ns, name)
                spec = lease.get("spec", {})
                holder, renew = spec.get("holderIdentity"), float(spec.get("renewTime", 0))
                if holder == self.identity or now - renew > LEASE_DURATION:
                    spec.update({"holderIdentity": self.identity, "renewTime": now,
                                 "leaseDurationSeconds": LEASE_DURATION})
                    lease["spec"] = spec
                    self.store.update("leases", lease)
                    self._set_leader(True)
                else:
                    self._set_leader(False)
            except ApiError as e:
                if e.code == 404:
                    try:
                        self.store.create("leases", {
                            "metadata": {"name": name, "namespace": ns},
                            "spec": {"holderIdentity": self.identity, "renewTime": now,
                                     "leaseDurationSeconds": LEASE_DURATION}}, ns)
                        self._set_leader(True)
                    except ApiError:
                        self._set_leader(False)
                else:
                    self._set_leader(False)   # lost a write race: retry
            self._stop.wait(RETRY_PERIOD if not self._leader.is_set() else RENEW_DEADLINE / 2)

    def _set_leader(self, v: bool) -> None:
        if v and not self._leader.is_set():
            log.info("%s became leader", self.identity)
            self._leader.set()
        elif not v and self._leader.is_set():
            log.error("leader election lost")
            self._leader.clear()
        self.metrics.is_leader.set(1 if v else 0)

    def is_leader(self) -> bool:
        return self._leader.is_set() or not self.opt.leader_elect

    # ------------------------------------------------------------------ http
    def _serve(self, addr: str, routes) -> Optional[ThreadingHTTPServer]:
        if not addr:
            return None
        host, _, port = addr.rpartition(":")

        class H(BaseHTTPRequestHandler):
            def do_GET(self):  # noqa: N802
                fn = routes.get(self.path.split("?")[0])
                if fn is None:
                    self.send_response(404)
                    self.end_headers()
                    return
                try:
                    ctype, body = fn(pprof.parse_query(self.path))
                except TypeError:
                    ctype, body = fn()
                self.send_response(200)
                self.send_header("Content-Type", ctype)
                self.send_header("Content-Length", str(len(body)))
                self.end_headers()
                self.wfile.write(body)

            def log_message(self, *a):
                pass

        srv = ThreadingHTTPServer((host or "0.0.0.0", int(port)), H)
        t = threading.Thread(target=srv.serve_forever, daemon=True, name=f"http-{addr}")
        t.start()
        self._servers.append(srv)
        return srv

    # ------------------------------------------------------------------ lifecycle
    def check_crd_exists(self) -> bool:
        """Legacy start-up gate (server.go:124-127,232-251)."""
        try:
            self.store.get("customresourcedefinitions", None, "tfjobs.kubeflow.org")
            return True
        except ApiError:
            return False

    def start(self) -> "Manager":
        if self.opt.check_crd and not self.check_crd_exists():
            raise RuntimeError("CRD tfjobs.kubeflow.org doesn't exist. Exiting")
        self.store.add_listener(self._on_event)
        # initial list (informer cache sync)
        for kind, r in self.reconcilers.items():
            for job in self.store.list(r.info.plural, self.opt.namespace or None):
                self.queues[kind].add(c.key_of(job))
        for kind in self.reconcilers:
            for i in range(max(1, self.opt.threadiness)):
                t = threading.Thread(target=self._worker, args=(kind,), daemon=True,
                                     name=f"{kind}-worker-{i}")
                t.start()
                self._threads.append(t)
        t = threading.Thread(target=self._resync, daemon=True, name="resync")
        t.start()
        self._threads.append(t)
        if self.opt.leader_elect:
            t = threading.Thread(target=self._elect, daemon=True, name="leader-election")
            t.start()
            self._threads.append(t)
        else:
            self.metrics.is_leader.set(1)
        self._serve(self.opt.metrics_bind_address, {
            "/metrics": lambda: ("text/plain; version=0.0.4", self.metrics.render()),
            # net/http/pprof shares the legacy binary's metrics mux (cmd/tf-operator.v1/main.go:21,
            # 39-50): goroutine / profile / heap / threadcreate / cmdline (operator/pprof.py)
            **pprof.routes()})
        self._serve(self.opt.health_probe_bind_address, {
            "/healthz": lambda: ("text/plain", b"ok"), "/readyz": lambda: ("text/plain", b"ok")})
        self.started = True
        return self

    def stop(self) -> None:
        self._stop.set()
        self.store.remove_listener(self._on_event)
        for q in self.queues.values():
            q.shut_down()
        for s in self._servers:
            s.shutdown()
            s.server_close()
        for t in self._threads:
            t.join(timeout=2)
        self.started = False

    def wait_idle(self, timeout: float = 5.0) -> bool:
        """Test helper: wait until all queues are drained and nothing is processing."""
        end = time.time() + timeout
        while time.time() < end:
            if all(len(q) == 0 and not q._processing and not q._dirty for q in self.queues.values()):
                return True
            time.sleep(0.005)
        return False
