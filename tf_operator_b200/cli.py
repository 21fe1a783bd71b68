"""``tfopctl``: the kubectl verbs the reference's docs use (docs/quick-start-v1.md: ``kubectl
apply -f``, ``get tfjobs``, ``describe``, ``logs``, ``delete``), against the operator's REST
façade on one box.

    scripts/setup-training-operator.sh 8443          # starts API server + operator + launcher
    export TFOP_APISERVER=http://127.0.0.1:8443
    python -m tf_operator_b200.cli apply -f examples/dist_mnist/tf_job_mnist.yaml
    python -m tf_operator_b200.cli get tfjobs
    python -m tf_operator_b200.cli describe tfjob dist-mnist
    python -m tf_operator_b200.cli logs dist-mnist-worker-0 [-f]
    python -m tf_operator_b200.cli wait tfjob dist-mnist --for Succeeded --timeout 300
    python -m tf_operator_b200.cli delete tfjob dist-mnist

Resource names: tfjob(s), pytorchjob(s), mxjob(s), xgboostjob(s), pod(s), service(s), event(s),
podgroup(s).
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time
from typing import Any, Dict, List, Optional

import yaml

from .api import KINDS, module_for
from .sdk.transport import HttpTransport
from .store.memory import ApiError

_CORE = {"pod": "pods", "pods": "pods", "service": "services", "services": "services", "svc": "services",
         "event": "events", "events": "events", "podgroup": "podgroups", "podgroups": "podgroups"}


def plural_of(resource: str) -> str:
    r = resource.lower()
    if r in _CORE:
        return _CORE[r]
    return module_for(r).INFO.plural


def _age(ts: Optional[str]) -> str:
    if not ts:
        return "<none>"
    from .api.common import parse_time
    s = max(0, int(time.time() - parse_time(ts)))
    for unit, n in (("d", 86400), ("h", 3600), ("m", 60)):
        if s >= n:
            return f"{s // n}{unit}"
    return f"{s}s"


def _state(obj: Dict[str, Any]) -> str:
    st = obj.get("status") or {}
    if "phase" in st:
        return st["phase"]
    conds = st.get("conditions") or []
    return conds[-1]["type"] if conds else ""


def _table(rows: List[List[str]]) -> str:
    widths = [max(len(r[i]) for r in rows) for i in range(len(rows[0]))]
    return "\n".join("   ".join(c.ljust(w) for c, w in zip(r, widths)).rstrip() for r in rows)


def cmd_apply(t: HttpTransport, args) -> int:
    docs = [d for d in yaml.safe_load_all(sys.stdin if args.filename == "-" else open(args.filename)) if d]
    for obj in docs:
        kind = obj.get("kind", "")
        if kind not in KINDS:
            print(f"error: unsupported kind {kind!r} (one of {sorted(KINDS)})", file=sys.stderr)
            return 1
        plural = KINDS[kind].INFO.plural
        ns = obj.get("metadata", {}).get("namespace") or args.namespace
        name = obj["metadata"]["name"]
        try:
            t.get(plural, ns, name)
            t.patch(plural, ns, name, obj)
            print(f"{plural[:-1]}.kubeflow.org/{name} configured")
        except ApiError as e:
            if e.code != 404:
                raise
            t.create(plural, ns, obj)
            print(f"{plural[:-1]}.kubeflow.org/{name} created")
    return 0


def cmd_get(t: HttpTransport, args) -> int:
    plural = plural_of(args.resource)
    if args.name:
        items = [t.get(plural, args.namespace, args.name)]
    else:
        items = t.list(plural, args.namespace, args.selector)
    if args.output == "json":
        print(json.dumps(items[0] if args.name else {"items": items}, indent=2))
    elif args.output == "yaml":
        print(yaml.safe_dump(items[0] if args.name else {"items": items}, sort_keys=False))
    elif args.output == "name":
        for o in items:
            print(f"{plural[:-1]}/{o['metadata']['name']}")
    else:
        if not items:
            print(f"No resources found in {args.namespace} namespace.")
            return 0
        if plural == "events":
            rows = [["LAST SEEN", "TYPE", "REASON", "OBJECT", "MESSAGE"]]
            for o in items:
                io = o.get("involvedObject", {})
                rows.append([_age(o.get("lastTimestamp")), o.get("type", ""), o.get("reason", ""),
                             f"{io.get('kind', '').lower()}/{io.get('name', '')}", o.get("message", "")])
        else:
            rows = [["NAME", "STATE", "AGE"]]
            for o in items:
                rows.append([o["metadata"]["name"], _state(o), _age(o["metadata"].get("creationTimestamp"))])
        print(_table(rows))
    return 0


def cmd_describe(t: HttpTransport, args) -> int:
    plural = plural_of(args.resource)
    obj = t.get(plural, args.namespace, args.name)
    md = obj["metadata"]
    print(f"Name:         {md['name']}\nNamespace:    {md.get('namespace', '')}")
    print(f"Kind:         {obj.get('kind', '')}\nUID:          {md.get('uid', '')}")
    if md.get("labels"):
        print("Labels:       " + ", ".join(f"{k}={v}" for k, v in md["labels"].items()))
    print("Spec:\n" + "\n".join("  " + l for l in yaml.safe_dump(obj.get("spec", {}), sort_keys=False).splitlines()))
    st = obj.get("status") or {}
    if st.get("conditions"):
        rows = [["  Type", "Status", "Reason", "Message"]]
        for c in st["conditions"]:
            rows.append(["  " + c.get("type", ""), c.get("status", ""), c.get("reason", ""), c.get("message", "")])
        print("Conditions:\n" + _table(rows))
    if st.get("replicaStatuses"):
        print("Replica Statuses:")
        for rt, rs in st["replicaStatuses"].items():
            print(f"  {rt}: " + ", ".join(f"{k}={v}" for k, v in rs.items()))
    evs = [e for e in t.list("events", args.namespace)
           if e.get("involvedObject", {}).get("uid") == md.get("uid")]
    if evs:
        rows = [["  Type", "Reason", "Age", "Message"]]
        for e in evs[-20:]:
            rows.append(["  " + e.get("type", ""), e.get("reason", ""), _age(e.get("lastTimestamp")),
                         e.get("message", "")])
        print("Events:\n" + _table(rows))
    return 0


def cmd_logs(t: HttpTransport, args) -> int:
    if args.follow:
        for line in t.pod_log_stream(args.namespace, args.pod, args.container):
            print(line, flush=True)
    else:
        sys.stdout.write(t.pod_log(args.namespace, args.pod, args.container))
    return 0


def cmd_delete(t: HttpTransport, args) -> int:
    plural = plural_of(args.resource)
    t.delete(plural, args.namespace, args.name)
    print(f'{plural[:-1]} "{args.name}" deleted')
    return 0


def cmd_wait(t: HttpTransport, args) -> int:
    plural = plural_of(args.resource)
    wanted = {w.strip() for w in getattr(args, "for").split(",")}
    end = time.time() + args.timeout
    while time.time() < end:
        obj = t.get(plural, args.namespace, args.name)
        conds = (obj.get("status") or {}).get("conditions") or []
        hit = [c["type"] for c in conds if c.get("type") in wanted and c.get("status") == "True"]
        if hit or _state(obj) in wanted:
            print(f"{plural[:-1]}/{args.name} condition met: {hit[0] if hit else _state(obj)}")
            return 0
        time.sleep(args.interval)
    print(f"error: timed out waiting for {sorted(wanted)} on {plural[:-1]}/{args.name}", file=sys.stderr)
    return 1


def build_parser() -> argparse.ArgumentParser:
    ap = argparse.ArgumentParser(prog="tfopctl", description="kubectl-style client for tf_operator_b200")
    ap.add_argument("--server", default=os.environ.get("TFOP_APISERVER", "http://127.0.0.1:8443"))
    ap.add_argument("-n", "--namespace", default="default")
    sub = ap.add_subparsers(dest="cmd", required=True)
    p = sub.add_parser("apply")
    p.add_argument("-f", "--filename", required=True)
    p = sub.add_parser("get")
    p.add_argument("resource")
    p.add_argument("name", nargs="?")
    p.add_argument("-o", "--output", default="", choices=["", "json", "yaml", "name"])
    p.add_argument("-l", "--selector", default=None)
    p = sub.add_parser("describe")
    p.add_argument("resource")
    p.add_argument("name")
    p = sub.add_parser("logs")
    p.add_argument("pod")
    p.add_argument("-c", "--container", default=None)
    p.add_argument("-f", "--follow", action="store_true")
    p = sub.add_parser("delete")
    p.add_argument("resource")
    p.add_argument("name")
    p = sub.add_parser("wait")
    p.add_argument("resource")
    p.add_argument("name")
    p.add_argument("--for", default="Succeeded,Failed")
    p.add_argument("--timeout", type=float, default=600)
    p.add_argument("--interval", type=float, default=0.2)
    return ap


def main(argv: Optional[List[str]] = None) -> int:
    args = build_parser().parse_args(argv)
    t = HttpTransport(args.server)
    try:
        return {"apply": cmd_apply, "get": cmd_get, "describe": cmd_describe, "logs": cmd_logs,
                "delete": cmd_delete, "wait": cmd_wait}[args.cmd](t, args)
    except ApiError as e:
        print(f"Error from server ({e.reason}): {e.message}", file=sys.stderr)
        return 1
    except KeyError as e:
        print(f"error: {e}", file=sys.stderr)
        return 1


if __name__ == "__main__":
    sys.exit(main())
