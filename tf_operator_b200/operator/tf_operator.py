"""Legacy ``tf-operator.v1`` entry point: TFJob only, client-go era flags
(cmd/tf-operator.v1/main.go:52-68, app/options/options.go:53-83, app/server.go:72-195).

    python -m tf_operator_b200.operator.tf_operator --threadiness 2 --monitoring-port 8443 \
        --enable-gang-scheduling --namespace default --qps 5 --burst 10

Differences of this binary from training-operator that are preserved: TFJob only; rate-limited
work queue with ``--threadiness`` workers; JSON log format by default; /metrics (and nothing
else) on ``--monitoring-port``; the CRD existence check at start-up; invalid specs are marked
Failed with reason InvalidTFJobSpec; leader election is always on; the client write throttle.
"""
from __future__ import annotations

import argparse
import json
import logging
import os
import signal
import sys
import threading

from .. import version
from ..api.common import ENV_KUBEFLOW_NAMESPACE
from ..cluster import LocalCluster
from ..launcher.kubelet import KubeletOptions
from .manager import ManagerOptions


class _JsonFormatter(logging.Formatter):
    """logrus JSON output with the `filename` hook (main.go:32-37,58-61)."""

    def format(self, record):
        return json.dumps({"level": record.levelname.lower(), "msg": record.getMessage(),
                           "time": self.formatTime(record, "%Y-%m-%dT%H:%M:%SZ"),
                           "filename": f"{record.filename}:{record.lineno}"})


def build_parser() -> argparse.ArgumentParser:
    p = argparse.ArgumentParser(prog="tf-operator.v1")
    p.add_argument("--master", default="", help="The url of the Kubernetes API server (the REST "
                   "façade of another process); empty = host one in this process")
    p.add_argument("--namespace", default=os.environ.get(ENV_KUBEFLOW_NAMESPACE, ""),
                   help="The namespace to monitor tfjobs. If unset, it monitors all namespaces.")
    p.add_argument("--threadiness", type=int, default=1, help="How many threads to process the main logic")
    p.add_argument("--version", action="store_true", help="Show version and quit")
    p.add_argument("--json-log-format", type=lambda s: s.lower() != "false", default=True,
                   help="Set true to use json style log format. Set false to use plaintext style log format")
    p.add_argument("--enable-gang-scheduling", action="store_true",
                   help="Set true to enable gang scheduling")
    p.add_argument("--gang-scheduler-name", default="volcano", help="The scheduler to gang-schedule tfjobs")
    p.add_argument("--monitoring-port", type=int, default=8443,
                   help="Endpoint port for displaying monitoring metrics. It can be set to \"0\" to disable")
    p.add_argument("--resyc-period", type=float, default=12 * 3600.0,
                   help="Resync interval of the tf-operator (seconds)")   # [sic] options.go:79
    p.add_argument("--qps", type=float, default=5, help="QPS indicates the maximum QPS to the master from this client.")
    p.add_argument("--burst", type=int, default=10, help="Maximum burst for throttle.")
    p.add_argument("--apiserver-port", type=int, default=8001)
    return p


def options_from_args(args) -> ManagerOptions:
    return ManagerOptions(
        enabled_schemes=["TFJob"], namespace=args.namespace, threadiness=args.threadiness,
        enable_gang_scheduling=args.enable_gang_scheduling,
        gang_scheduler_name=args.gang_scheduler_name, leader_elect=True,
        leader_election_id="tf-operator",
        metrics_bind_address=f":{args.monitoring_port}" if args.monitoring_port else "",
        fail_invalid_spec=True, check_crd=True, qps=args.qps, burst=args.burst)


def main(argv=None) -> int:
    args = build_parser().parse_args(argv)
    if args.version:
        version.print_version_and_exit()
    handler = logging.StreamHandler()
    if args.json_log_format:
        handler.setFormatter(_JsonFormatter())
    logging.basicConfig(level=logging.INFO, handlers=[handler])
    logging.getLogger(__name__).info("%s", version.info())
    cluster = LocalCluster(manager_options=options_from_args(args),
                           serve_http=args.apiserver_port != 0, http_port=args.apiserver_port,
                           kubelet_options=KubeletOptions())
    stop = threading.Event()
    for sig in (signal.SIGINT, signal.SIGTERM):
        signal.signal(sig, lambda *_: stop.set())
    stop.wait()
    cluster.stop()
    return 0


if __name__ == "__main__":
    sys.exit(main())
