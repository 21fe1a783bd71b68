"""Name -> address mapping for one box ("cluster DNS").

In the reference every replica gets a headless Service whose DNS name
``<job>-<type>-<index>[.<ns>.svc[.<domain>]]`` resolves to the pod
(pkg/controller.v1/tensorflow/tensorflow.go:153-166). On one box there is no DNS, but the
whole 127.0.0.0/8 block is loopback on Linux, so every pod gets its own loopback IP derived
from (cluster id, namespace, name). All replicas can therefore keep the SAME well-known port
(2222, 23456, ...) without colliding, and any process can compute a peer's address from the name
alone — no lookup service, nothing to race with pod start-up order. The cluster id (one per
LocalCluster, exported to replicas as ``TFOP_CLUSTER_ID``) keeps two clusters on one host - two
operators, or parallel test sessions creating same-named jobs - on different addresses.
"""
from __future__ import annotations

import hashlib
import os
from typing import Tuple

ENV_NAMESPACE = "TFOP_POD_NAMESPACE"
ENV_CLUSTER_ID = "TFOP_CLUSTER_ID"


def pod_ip(namespace: str, name: str, cluster_id: str | None = None) -> str:
    if cluster_id is None:
        cluster_id = os.environ.get(ENV_CLUSTER_ID, "")
    key = f"{cluster_id}|{namespace}/{name}" if cluster_id else f"{namespace}/{name}"
    h = hashlib.sha1(key.encode()).digest()
    a, b, c_ = h[0], h[1], h[2]
    if a == 0 and b == 0:
        b = 1                     # stay out of 127.0.0.x
    c_ = min(max(c_, 1), 254)
    return f"127.{a}.{b}.{c_}"


def split_host(host: str, default_ns: str) -> Tuple[str, str]:
    """'job-worker-0.ns.svc.cluster.local' -> ('ns', 'job-worker-0'); bare names use default_ns."""
    parts = host.split(".")
    if len(parts) >= 3 and parts[2] == "svc":
        return parts[1], parts[0]
    if len(parts) == 2:
        return parts[1], parts[0]
    return default_ns, parts[0]


def resolve(host: str, namespace: str | None = None) -> str:
    """Resolve a replica/service name to its loopback IP. IPs and 'localhost' pass through."""
    if host in ("localhost", "") or host.replace(".", "").isdigit():
        return host or "127.0.0.1"
    ns = namespace or os.environ.get(ENV_NAMESPACE, "default")
    ns, name = split_host(host, ns)
    return pod_ip(ns, name)


def resolve_endpoint(endpoint: str, namespace: str | None = None) -> str:
    """'host:port' -> 'ip:port'."""
    host, sep, port = endpoint.rpartition(":")
    if not sep:
        return resolve(endpoint, namespace)
    return f"{resolve(host, namespace)}:{port}"
