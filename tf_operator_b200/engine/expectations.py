"""Controller expectations: per-key counters of creations/deletions the controller has
issued but not yet observed, so a reconcile against a stale cache does not double-create.

Semantics of k8s ``ControllerExpectations`` as used through kubeflow/common
(pinned by pkg/controller.v1/tensorflow/pod_test.go:109-228;
glue in pkg/common/util/reconciler.go:23-62).
"""
from __future__ import annotations

import threading
import time
from typing import Dict, Iterable

EXPECTATION_TIMEOUT_S = 5 * 60


def gen_expectation_pods_key(job_key: str, rtype: str) -> str:
    return f"{job_key}/{rtype.lower()}/pods"


def gen_expectation_services_key(job_key: str, rtype: str) -> str:
    return f"{job_key}/{rtype.lower()}/services"


class _Entry:
    __slots__ = ("add", "dele", "ts")

    def __init__(self, add: int, dele: int):
        self.add, self.dele, self.ts = add, dele, time.monotonic()

    def fulfilled(self) -> bool:
        return self.add <= 0 and self.dele <= 0

    def expired(self) -> bool:
        return time.monotonic() - self.ts > EXPECTATION_TIMEOUT_S


class ControllerExpectations:
    def __init__(self):
        self._lock = threading.Lock()
        self._store: Dict[str, _Entry] = {}

    def get(self, key: str):
        with self._lock:
            e = self._store.get(key)
            return None if e is None else (e.add, e.dele)

    def satisfied_expectations(self, key: str) -> bool:
        """True when fulfilled, expired, or never set (a new controller must sync)."""
        with self._lock:
            e = self._store.get(key)
            if e is None:
                return True
            return e.fulfilled() or e.expired()

    def set_expectations(self, key: str, add: int, dele: int) -> None:
        with self._lock:
            self._store[key] = _Entry(add, dele)

    def expect_creations(self, key: str, n: int) -> None:
        self.raise_expectations(key, n, 0)

    def expect_deletions(self, key: str, n: int) -> None:
        self.raise_expectations(key, 0, n)

    def raise_expectations(self, key: str, add: int, dele: int) -> None:
        with self._lock:
            e = self._store.get(key)
            if e is None:
                self._store[key] = _Entry(add, dele)
            else:
                e.add += add
                e.dele += dele

    def lower_expectations(self, key: str, add: int, dele: int) -> None:
        with self._lock:
            e = self._store.get(key)
            if e is not None:
                e.add -= add
                e.dele -= dele

    def creation_observed(self, key: str) -> None:
        self.lower_expectations(key, 1, 0)

    def deletion_observed(self, key: str) -> None:
        self.lower_expectations(key, 0, 1)

    def delete_expectations(self, key: str) -> None:
        with self._lock:
            self._store.pop(key, None)


def satisfied_expectations(exp: ControllerExpectations, job_key: str,
                           replica_types: Iterable[str]) -> bool:
    """OR over replica types of pods-key / services-key satisfied
    (pkg/common/util/reconciler.go:23-35)."""
    ok = False
    for rt in replica_types:
        ok = ok or exp.satisfied_expectations(gen_expectation_pods_key(job_key, rt))
        ok = ok or exp.satisfied_expectations(gen_expectation_services_key(job_key, rt))
    return ok
