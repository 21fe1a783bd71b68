"""Rate-limited, de-duplicating work queue with delayed adds.

client-go's ``workqueue.RateLimitingInterface`` as the legacy controller uses it
(pkg/controller.v1/tensorflow/controller.go:193-286): one in-flight item per key,
``AddRateLimited`` with per-item exponential back-off, ``AddAfter`` for the
activeDeadlineSeconds / TTL timers, ``NumRequeues`` for the back-off limit.
The new manager's ``FakeWorkQueue`` (pkg/common/util/fake_workqueue.go:5-36) turns all
of this into no-ops, which silently disables those timers;
we always run the real one.
"""
from __future__ import annotations

import heapq
import threading
import time
from typing import Dict, List, Optional, Set, Tuple


class RateLimitingQueue:
    def __init__(self, base_delay: float = 0.005, max_delay: float = 1000.0, name: str = ""):
        self.name = name
        self._cv = threading.Condition()
        self._queue: List[str] = []
        self._dirty: Set[str] = set()
        self._processing: Set[str] = set()
        self._delayed: List[Tuple[float, int, str]] = []
        self._seq = 0
        self._failures: Dict[str, int] = {}
        self._base, self._max = base_delay, max_delay
        self._shutdown = False
        self._timer = threading.Thread(target=self._timer_loop, daemon=True,
                                       name=f"workqueue-timer-{name}")
        self._timer.start()

    # ---- basic queue -------------------------------------------------------
    def add(self, key: str) -> None:
        with self._cv:
            if self._shutdown or key in self._dirty:
                return
            self._dirty.add(key)
            if key in self._processing:
                return
            self._queue.append(key)
            self._cv.notify_all()

    def get(self, timeout: Optional[float] = None) -> Optional[str]:
        """Block for the next key; None on shutdown/timeout. Pair with done(key)."""
        end = None if timeout is None else time.monotonic() + timeout
        with self._cv:
            while not self._queue and not self._shutdown:
                remaining = None if end is None else end - time.monotonic()
                if remaining is not None and remaining <= 0:
                    return None
                self._cv.wait(remaining)
            if not self._queue:
                return None
            key = self._queue.pop(0)
            self._processing.add(key)
            self._dirty.discard(key)
            return key

    def done(self, key: str) -> None:
        with self._cv:
            self._processing.discard(key)
            if key in self._dirty:
                self._queue.append(key)
                self._cv.notify_all()

    def __len__(self) -> int:
        with self._cv:
            return len(self._queue)

    # ---- delays / rate limiting ----------------------------------------------
    def add_after(self, key: str, delay: float) -> None:
        if delay <= 0:
            return self.add(key)
        with self._cv:
            if self._shutdown:
                return
            self._seq += 1
            heapq.heappush(self._delayed, (time.monotonic() + delay, self._seq, key))
            self._cv.notify_all()

    def add_rate_limited(self, key: str) -> None:
        with self._cv:
            n = self._failures.get(key, 0)
            self._failures[key] = n + 1
        self.add_after(key, min(self._base * (2 ** n), self._max))

    def forget(self, key: str) -> None:
        with self._cv:
            self._failures.pop(key, None)

    def num_requeues(self, key: str) -> int:
        with self._cv:
            return self._failures.get(key, 0)

    def _timer_loop(self):
        while True:
            with self._cv:
                if self._shutdown:
                    return
                now = time.monotonic()
                due = []
                while self._delayed and self._delayed[0][0] <= now:
                    due.append(heapq.heappop(self._delayed)[2])
                wait = (self._delayed[0][0] - now) if self._delayed else 0.5
            for k in due:
                self.add(k)
            with self._cv:
                if self._shutdown:
                    return
                if not due:
                    self._cv.wait(min(wait, 0.5))

    def shut_down(self) -> None:
        with self._cv:
            self._shutdown = True
            self._cv.notify_all()


class FakeWorkQueue:
    """No-op queue with the same surface (pkg/common/util/fake_workqueue.go:5-36);
    used by unit tests that call reconcile directly."""

    def add(self, key): pass
    def add_after(self, key, delay): pass
    def add_rate_limited(self, key): pass
    def forget(self, key): pass
    def num_requeues(self, key): return 0
    def done(self, key): pass
    def get(self, timeout=None): return None
    def shut_down(self): pass
    def __len__(self): return 0
