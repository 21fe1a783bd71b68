"""``/debug/pprof/*`` for the operator process - the endpoints Go's net/http/pprof gives the
legacy binary on its monitoring port (cmd/tf-operator.v1/main.go:21,39-50), with Python's own
instruments behind them:

  /debug/pprof/               index
  /debug/pprof/goroutine      every thread's current stack (``?debug=2`` accepted)
  /debug/pprof/profile        sampling CPU profile, ``?seconds=N`` (default 5, at 100 Hz): folded
                              stacks ``frame;frame;frame count`` - the flamegraph.pl / speedscope
                              input format
  /debug/pprof/heap           tracemalloc top allocation sites (tracing starts on first request)
                              plus gc generation counts
  /debug/pprof/threadcreate   live threads with daemon flag
  /debug/pprof/cmdline        the process command line, NUL separated as Go does
"""
from __future__ import annotations

import gc
import sys
import threading
import time
import traceback
import tracemalloc
from collections import Counter
from urllib.parse import parse_qs

INDEX = b"""/debug/pprof/
profiles:
goroutine
heap
profile
threadcreate
cmdline
"""


def goroutine(_q=None) -> bytes:
    names = {t.ident: t.name for t in threading.enumerate()}
    out = []
    for ident, frame in sys._current_frames().items():
        out.append(f"thread {names.get(ident, '?')} ({ident}):\n"
                   + "".join(traceback.format_stack(frame)))
    return "\n".join(out).encode()


def cpu_profile(q=None) -> bytes:
    secs = min(float((q or {}).get("seconds", ["5"])[0]), 60.0)
    me = threading.get_ident()
    names = {t.ident: t.name for t in threading.enumerate()}
    folded: Counter = Counter()
    end = time.monotonic() + secs
    while time.monotonic() < end:
        for ident, frame in sys._current_frames().items():
            if ident == me:
                continue
            stack = []
            while frame is not None:
                code = frame.f_code
                stack.append(f"{code.co_name} ({code.co_filename.rsplit('/', 1)[-1]}:{frame.f_lineno})")
                frame = frame.f_back
            folded[";".join([names.get(ident, "?")] + stack[::-1])] += 1
        time.sleep(0.01)
    return "".join(f"{k} {v}\n" for k, v in folded.most_common()).encode()


def heap(_q=None) -> bytes:
    if not tracemalloc.is_tracing():
        tracemalloc.start(8)
        return b"tracemalloc started; request again for allocation sites\n"
    snap = tracemalloc.take_snapshot()
    lines = [f"gc counts {gc.get_count()}  objects {len(gc.get_objects())}"]
    cur, peak = tracemalloc.get_traced_memory()
    lines.append(f"traced current {cur} B  peak {peak} B")
    for st in snap.statistics("lineno")[:50]:
        lines.append(str(st))
    return ("\n".join(lines) + "\n").encode()


def threadcreate(_q=None) -> bytes:
    return "".join(f"{t.name}\tdaemon={t.daemon}\talive={t.is_alive()}\n"
                   for t in threading.enumerate()).encode()


def cmdline(_q=None) -> bytes:
    return "\x00".join(sys.argv).encode()


def routes() -> dict:
    """path -> callable(query dict) -> (content type, body)"""
    text = "text/plain; charset=utf-8"
    return {
        "/debug/pprof/": lambda q: (text, INDEX),
        "/debug/pprof/goroutine": lambda q: (text, goroutine(q)),
        "/debug/pprof/profile": lambda q: (text, cpu_profile(q)),
        "/debug/pprof/heap": lambda q: (text, heap(q)),
        "/debug/pprof/threadcreate": lambda q: (text, threadcreate(q)),
        "/debug/pprof/cmdline": lambda q: (text, cmdline(q)),
    }


def parse_query(path: str) -> dict:
    return parse_qs(path.split("?", 1)[1]) if "?" in path else {}
