"""Token-bucket client throttle: client-go's flowcontrol limiter, which caps the reference
operator's API writes at QPS 5 / burst 10 by default
(cmd/tf-operator.v1/app/options/options.go:81-82, server.go:110-111)."""
from __future__ import annotations

import threading
import time


class TokenBucket:
    def __init__(self, qps: float, burst: int):
        self.qps, self.burst = float(qps), int(burst)
        self._tokens = float(burst)
        self._last = time.monotonic()
        self._lock = threading.Lock()

    def accept(self) -> float:
        """Block until a token is available; returns the time waited."""
        waited = 0.0
        while True:
            with self._lock:
                now = time.monotonic()
                self._tokens = min(self.burst, self._tokens + (now - self._last) * self.qps)
                self._last = now
                if self._tokens >= 1.0:
                    self._tokens -= 1.0
                    return waited
                need = (1.0 - self._tokens) / self.qps
            time.sleep(need)
            waited += need


class ThrottledControl:
    """Wraps a pod/service control so every write first takes a token."""

    def __init__(self, inner, bucket: TokenBucket):
        self._inner, self._bucket = inner, bucket

    def __getattr__(self, name):
        attr = getattr(self._inner, name)
        if callable(attr) and (name.startswith("create_") or name.startswith("delete_")
                               or name.startswith("patch_")):
            def wrapped(*a, **kw):
                self._bucket.accept()
                return attr(*a, **kw)
            return wrapped
        return attr
