"""``training-operator`` entry point: all four kinds, controller-runtime style flags
(cmd/training-operator.v1/main.go:58-124).

    python -m tf_operator_b200.operator.training_operator \
        --metrics-bind-address :8080 --health-probe-bind-address :8081 \
        [--leader-elect] [--enable-scheme tfjob --enable-scheme pytorchjob ...] \
        [--apiserver-port 8001] [--gpus-per-replica 0]

On one box the process also hosts the API server façade and the launcher (there is no
external kube-apiserver / kubelet to connect to).
"""
from __future__ import annotations

import argparse
import logging
import signal
import sys
import threading

from ..cluster import LocalCluster
from ..launcher.kubelet import KubeletOptions
from .manager import EnabledSchemes, ManagerOptions


def build_parser() -> argparse.ArgumentParser:
    p = argparse.ArgumentParser(prog="training-operator")
    p.add_argument("--metrics-bind-address", default=":8080",
                   help="The address the metric endpoint binds to.")
    p.add_argument("--health-probe-bind-address", default=":8081",
                   help="The address the probe endpoint binds to.")
    p.add_argument("--leader-elect", action="store_true",
                   help="Enable leader election for controller manager. Enabling this will ensure "
                        "there is only one active controller manager.")
    p.add_argument("--enable-scheme", action="append", default=[],
                   help="Enable scheme(s) as --enable-scheme=tfjob --enable-scheme=pytorchjob, "
                        "case insensitive. Now supporting TFJob, PyTorchJob, MXNetJob, XGBoostJob. "
                        "By default, all supported schemes will be enabled.")
    p.add_argument("--zap-devel", action="store_true", default=True, help="development-mode logging")
    p.add_argument("--zap-log-level", default="info")
    # single-box additions
    p.add_argument("--apiserver-port", type=int, default=8001, help="REST façade port (0 = off)")
    p.add_argument("--gpus-per-replica", type=int, default=0,
                   help="GPUs for replicas that do not request nvidia.com/gpu")
    p.add_argument("--enable-gang-scheduling", action="store_true")
    return p


def options_from_args(args) -> ManagerOptions:
    schemes = EnabledSchemes()
    for s in args.enable_scheme:
        schemes.set({"mxnetjob": "MXJob"}.get(s.lower(), s))
    return ManagerOptions(enabled_schemes=list(schemes), leader_elect=args.leader_elect,
                          metrics_bind_address=args.metrics_bind_address,
                          health_probe_bind_address=args.health_probe_bind_address,
                          enable_gang_scheduling=args.enable_gang_scheduling)


def main(argv=None) -> int:
    args = build_parser().parse_args(argv)
    logging.basicConfig(level=getattr(logging, args.zap_log_level.upper(), logging.INFO),
                        format="%(asctime)s %(levelname)s %(name)s %(message)s")
    log = logging.getLogger("setup")
    try:
        opts = options_from_args(args)
    except ValueError as e:
        log.error("unable to parse --enable-scheme: %s", e)
        return 1
    cluster = LocalCluster(manager_options=opts, serve_http=args.apiserver_port != 0,
                           http_port=args.apiserver_port,
                           kubelet_options=KubeletOptions(gpus_per_replica_default=args.gpus_per_replica))
    log.info("starting manager: schemes=%s api=%s", cluster.manager.schemes,
             cluster.api.url if cluster.api else "in-process only")
    stop = threading.Event()
    for sig in (signal.SIGINT, signal.SIGTERM):
        signal.signal(sig, lambda *_: stop.set())
    stop.wait()
    cluster.stop()
    return 0


if __name__ == "__main__":
    sys.exit(main())
