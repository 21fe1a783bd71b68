"""Operator process: scheme registry, flags of both binaries, metrics / probes endpoints,
leader election, work queue, namespace scoping (cmd/*, pkg/controller.v1/register_controller.go)."""
import threading
import time
import urllib.request

import pytest

from tf_operator_b200.api import common as c
from tf_operator_b200.engine.workqueue import RateLimitingQueue
from tf_operator_b200.operator import tf_operator, training_operator
from tf_operator_b200.operator.manager import EnabledSchemes, Manager, ManagerOptions
from tf_operator_b200.store.memory import ObjectStore


def _tfjob(name, ns="default"):
    return {"apiVersion": "kubeflow.org/v1", "kind": "TFJob", "metadata": {"name": name, "namespace": ns},
            "spec": {"tfReplicaSpecs": {"Worker": {"replicas": 1, "template": {"spec": {"containers": [
                {"name": "tensorflow", "image": "x"}]}}}}}}


def test_enabled_schemes_case_insensitive_and_fill_all():
    s = EnabledSchemes()
    assert s.empty()
    s.set("tfjob"); s.set("PYTORCHJOB"); s.set("tfjob")
    assert list(s) == ["TFJob", "PyTorchJob"]
    with pytest.raises(ValueError):
        s.set("caffejob")
    s.fill_all()
    assert set(s) == {"TFJob", "PyTorchJob", "MXJob", "XGBoostJob"}


def test_training_operator_flags():
    args = training_operator.build_parser().parse_args(
        ["--enable-scheme", "tfjob", "--enable-scheme=XGBoostJob", "--leader-elect",
         "--metrics-bind-address", ":9090"])
    o = training_operator.options_from_args(args)
    assert o.enabled_schemes == ["TFJob", "XGBoostJob"] and o.leader_elect
    assert o.metrics_bind_address == ":9090" and o.health_probe_bind_address == ":8081"
    d = training_operator.build_parser().parse_args([])
    assert d.metrics_bind_address == ":8080" and not d.leader_elect and d.enable_scheme == []


def test_legacy_tf_operator_flags():
    d = tf_operator.build_parser().parse_args([])
    assert (d.threadiness, d.monitoring_port, d.qps, d.burst, d.gang_scheduler_name) == (1, 8443, 5, 10, "volcano")
    assert d.json_log_format is True and d.enable_gang_scheduling is False and d.resyc_period == 43200
    o = tf_operator.options_from_args(tf_operator.build_parser().parse_args(
        ["--namespace", "team-a", "--threadiness", "4", "--enable-gang-scheduling", "--monitoring-port", "0"]))
    assert o.enabled_schemes == ["TFJob"] and o.namespace == "team-a" and o.threadiness == 4
    assert o.enable_gang_scheduling and o.metrics_bind_address == "" and o.fail_invalid_spec and o.check_crd
    assert o.leader_elect and o.leader_election_id == "tf-operator" and o.qps == 5


def test_metrics_and_probe_endpoints_and_counters():
    store = ObjectStore()
    m = Manager(store, ManagerOptions(metrics_bind_address="127.0.0.1:0", health_probe_bind_address="127.0.0.1:0")).start()
    try:
        mport, hport = (s.server_address[1] for s in m._servers)
        assert urllib.request.urlopen(f"http://127.0.0.1:{hport}/healthz").read() == b"ok"
        assert urllib.request.urlopen(f"http://127.0.0.1:{hport}/readyz").read() == b"ok"
        store.create("tfjobs", _tfjob("m1"))
        assert m.wait_idle()
        for p in store.list("pods"):
            p["status"] = {"phase": "Succeeded", "containerStatuses": [
                {"name": "tensorflow", "state": {"terminated": {"exitCode": 0}}}]}
            store.update_status("pods", p)
        assert m.wait_idle()
        store.delete("tfjobs", "default", "m1")
        text = urllib.request.urlopen(f"http://127.0.0.1:{mport}/metrics").read().decode()
        assert 'tf_operator_jobs_created_total{job_namespace="default"} 1.0' in text
        assert 'tf_operator_jobs_successful_total{job_namespace="default"}' in text
        assert "tf_operator_is_leader 1.0" in text
    finally:
        m.stop()


def test_namespace_scoping():
    store = ObjectStore()
    m = Manager(store, ManagerOptions(enabled_schemes=["TFJob"], namespace="team-a")).start()
    try:
        store.create("tfjobs", _tfjob("in", "team-a"))
        store.create("tfjobs", _tfjob("out", "team-b"))
        assert m.wait_idle()
        time.sleep(0.1)
        assert [c.name_of(p) for p in store.list("pods")] == ["in-worker-0"]
    finally:
        m.stop()


def test_legacy_crd_check_refuses_to_start_without_crd():
    with pytest.raises(RuntimeError, match="tfjobs.kubeflow.org"):
        Manager(ObjectStore(), ManagerOptions(check_crd=True)).start()


def test_leader_election_only_one_active(monkeypatch):
    from tf_operator_b200.operator import manager as mod
    monkeypatch.setattr(mod, "RETRY_PERIOD", 0.05)
    monkeypatch.setattr(mod, "RENEW_DEADLINE", 0.1)
    monkeypatch.setattr(mod, "LEASE_DURATION", 0.4)
    store = ObjectStore()
    a = Manager(store, ManagerOptions(leader_elect=True, identity="a")).start()
    time.sleep(0.15)
    b = Manager(store, ManagerOptions(leader_elect=True, identity="b")).start()
    try:
        time.sleep(0.3)
        assert a.is_leader() and not b._leader.is_set()
        store.create("tfjobs", _tfjob("le"))
        time.sleep(0.3)
        assert len(store.list("pods")) == 1            # exactly one manager reconciled
        a.stop()
        end = time.time() + 3
        while time.time() < end and not b._leader.is_set():
            time.sleep(0.02)
        assert b._leader.is_set()                       # takes over after the lease expires
    finally:
        b.stop()
        if a.started:
            a.stop()


def test_rate_limiting_queue_dedup_delay_and_backoff():
    q = RateLimitingQueue(base_delay=0.01, max_delay=0.05)
    q.add("a"); q.add("a"); q.add("b")
    assert len(q) == 2
    k = q.get(timeout=1)
    q.add(k)                       # re-added while processing: queued again after done()
    assert len(q) == 1
    q.done(k)
    assert len(q) == 2
    assert {q.get(timeout=1), q.get(timeout=1)} == {"a", "b"}
    q.done("a"); q.done("b")
    t0 = time.time()
    q.add_after("later", 0.1)
    assert q.get(timeout=0.03) is None
    assert q.get(timeout=1) == "later" and time.time() - t0 >= 0.09
    q.done("later")
    for _ in range(3):
        q.add_rate_limited("x")
    assert q.num_requeues("x") == 3
    q.forget("x")
    assert q.num_requeues("x") == 0
    q.shut_down()
    assert q.get(timeout=0.1) is None


def test_active_deadline_fires_without_any_other_event():
    """the reference's new manager never re-queues for the deadline
    (FakeWorkQueue); ours does."""
    store = ObjectStore()
    m = Manager(store, ManagerOptions(enabled_schemes=["TFJob"])).start()
    try:
        job = _tfjob("dl")
        job["spec"]["runPolicy"] = {"activeDeadlineSeconds": 1}
        store.create("tfjobs", job)
        assert m.wait_idle()
        for p in store.list("pods"):
            p["status"] = {"phase": "Running"}
            store.update_status("pods", p)
        end = time.time() + 5
        failed = None
        while time.time() < end and not failed:
            st = store.get("tfjobs", "default", "dl").get("status") or {}
            failed = c.get_condition(st, "Failed") if st.get("conditions") else None
            time.sleep(0.05)
        assert failed and "deadline" in failed["message"]
        assert store.list("pods") == []
    finally:
        m.stop()


def test_pprof_endpoints_and_sync_duration_log(caplog):
    """The monitoring port serves the net/http/pprof set (cmd/tf-operator.v1/main.go:21,39-50)
    and every sync logs its wall time (controller.go:303-307 "Finished syncing tfjob %q (%v)")."""
    import logging
    store = ObjectStore()
    m = Manager(store, ManagerOptions(metrics_bind_address="127.0.0.1:0",
                                      health_probe_bind_address="127.0.0.1:0")).start()
    try:
        mport = m._servers[0].server_address[1]

        def get(path):
            return urllib.request.urlopen(f"http://127.0.0.1:{mport}{path}").read().decode()
        idx = get("/debug/pprof/")
        for name in ("goroutine", "heap", "profile", "threadcreate", "cmdline"):
            assert name in idx
        assert "TFJob-worker-0" in get("/debug/pprof/goroutine?debug=2")
        prof = get("/debug/pprof/profile?seconds=0.3")
        assert "_worker (manager.py:" in prof and prof.strip().split()[-1].isdigit()
        assert "tracemalloc" in get("/debug/pprof/heap") or "traced current" in get("/debug/pprof/heap")
        assert "traced current" in get("/debug/pprof/heap")
        assert "resync\tdaemon=True" in get("/debug/pprof/threadcreate")
        assert get("/debug/pprof/cmdline")
        with caplog.at_level(logging.INFO, logger="tf_operator_b200.operator"):
            store.create("tfjobs", _tfjob("sync-log"))
            assert m.wait_idle()
        msgs = [r.getMessage() for r in caplog.records if "Finished syncing" in r.getMessage()]
        assert msgs and "Finished syncing tfjob 'default/sync-log' (" in msgs[0] and msgs[0].endswith("ms)")
        text = get("/metrics")
        assert 'training_operator_reconcile_seconds_count{kind="TFJob"}' in text
    finally:
        m.stop()
        import tracemalloc
        tracemalloc.stop()


def test_two_workers_never_mix_up_the_pods_of_concurrent_tfjobs():
    """the TFJob reconciler is shared by `threadiness` worker threads; the pod list it
    remembers between reconcile_pods() and update_job_status() must be the calling thread's own.
    Job a's worker-0 has succeeded, job b's is still running: with the syncs of a and b forced to
    interleave, b must not be marked Succeeded on the strength of a's pods."""
    from tf_operator_b200.kinds.tf import TFJobReconciler
    store = ObjectStore()
    m = Manager(store, ManagerOptions(threadiness=2, enabled_schemes=["TFJob"],
                                      metrics_bind_address="", health_probe_bind_address="")).start()
    gate = threading.Barrier(2, timeout=2)
    orig = TFJobReconciler.update_job_status
    seen = []

    def slow_update(self, job, replicas, status):
        seen.append((threading.current_thread().name, c.name_of(job)))
        try:
            gate.wait()        # both threads have stored their pod lists before either reads one
        except threading.BrokenBarrierError:
            pass
        return orig(self, job, replicas, status)

    try:
        for name in ("a", "b"):
            job = _tfjob(name)
            job["spec"]["tfReplicaSpecs"]["Worker"]["replicas"] = 2
            store.create("tfjobs", job)
        assert m.wait_idle()
        TFJobReconciler.update_job_status = slow_update
        for p in store.list("pods"):
            done = p["metadata"]["name"] == "a-worker-0"
            p["status"] = {"phase": "Succeeded" if done else "Running", "containerStatuses": [
                {"name": "tensorflow",
                 "state": {"terminated": {"exitCode": 0}} if done else {"running": {}}}]}
            store.update_status("pods", p)
        assert m.wait_idle(timeout=15)
        time.sleep(0.3)
        assert m.wait_idle(timeout=15)
    finally:
        TFJobReconciler.update_job_status = orig
        m.stop()
    assert len({t for t, _ in seen}) == 2, seen        # both worker threads took part
    conds = {n: [cd["type"] for cd in store.get("tfjobs", "default", n)["status"]["conditions"]]
             for n in ("a", "b")}
    assert "Succeeded" in conds["a"], conds
    assert "Succeeded" not in conds["b"] and "Running" in conds["b"], conds
