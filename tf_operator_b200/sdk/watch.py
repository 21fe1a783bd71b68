"""``watch``: stream job events and print a NAME / STATE / TIME table until the named job
finishes (sdk/python/kubeflow/tfjob/api/tf_job_watch.py:23-59; table_logger and retrying are
not installed, so the table and the 20 x 1 s retry are done by hand)."""
from __future__ import annotations

import sys
import time


def _row(name, state, ts) -> str:
    return f"{name:<30.30} {state:<20.20} {ts:<30.30}"


def watch(client, name=None, namespace=None, timeout_seconds=600, out=None):
    out = out or sys.stdout
    out.write(_row("NAME", "STATE", "TIME") + "\n")
    last_err = None
    for _attempt in range(20):
        try:
            for _etype, job in client.transport.watch(client.plural, namespace, timeout_seconds):
                jname = job.get("metadata", {}).get("name", "")
                if name and jname != name:
                    continue
                conds = (job.get("status") or {}).get("conditions") or []
                if conds:
                    last = conds[-1]
                    out.write(_row(jname, last.get("type", ""), last.get("lastTransitionTime", "")) + "\n")
                    out.flush()
                    if name and last.get("type") in ("Succeeded", "Failed"):
                        return
                else:
                    out.write(_row(jname, "", "") + "\n")
            return
        except (OSError, RuntimeError) as e:   # transient: retry like @retrying.retry
            last_err = e
            time.sleep(1)
    if last_err is not None:
        raise last_err
