"""LocalCluster: API server + operator + launcher + rendezvous in one process — the
"single-box cluster" a TFJob / PyTorchJob / MXJob / XGBoostJob is submitted to.

    with LocalCluster() as cluster:
        client = PyTorchJobClient(cluster=cluster)
        client.create(job); client.wait_for_job("name")

Stands in for what the reference needs a Kubernetes cluster for (SURVEY.md layer map L0,
L3-L5): kube-apiserver (store/), the operator Deployment (operator/), kubelet + Volcano
(launcher/). ``serve_http=True`` also exposes the REST façade so out-of-process SDK clients
(TFOP_APISERVER=http://...) and the replicas' fault-injection proxy work.
"""
from __future__ import annotations

import itertools
import os
import threading
from typing import Optional

from .launcher.kubelet import KubeletOptions, LocalKubelet
from .launcher.rendezvous import RendezvousServer
from .launcher.topology import Topology
from .operator.manager import Manager, ManagerOptions
from .store.http import ApiServer
from .store.memory import ObjectStore

_default_lock = threading.Lock()
_cluster_seq = itertools.count(1)
_default: Optional["LocalCluster"] = None


class LocalCluster:
    def __init__(self, *, manager_options: Optional[ManagerOptions] = None,
                 kubelet_options: Optional[KubeletOptions] = None,
                 topology: Optional[Topology] = None, serve_http: bool = False,
                 http_port: int = 0, start: bool = True):
        self.store = ObjectStore()
        self.rendezvous = RendezvousServer()
        kopt = kubelet_options or KubeletOptions()
        if topology is not None:
            kopt.topology = topology
        if not kopt.cluster_id:
            # unique per cluster object AND per host process: pod addresses of two clusters differ
            kopt.cluster_id = f"{os.getpid():x}-{next(_cluster_seq)}"
        self.cluster_id = kopt.cluster_id
        kopt.rendezvous_addr = self.rendezvous.addr
        kopt.rendezvous_uds = self.rendezvous.uds_path
        self.kubelet = LocalKubelet(self.store, kopt)
        self.manager = Manager(self.store, manager_options or ManagerOptions())
        self.api: Optional[ApiServer] = ApiServer(self.store, self.kubelet, port=http_port) \
            if serve_http else None
        self._register_crds()
        self.running = False
        if start:
            self.start()

    def _register_crds(self) -> None:
        from .api import KINDS
        for mod in KINDS.values():
            i = mod.INFO
            self.store.create("customresourcedefinitions", {
                "apiVersion": "apiextensions.k8s.io/v1", "kind": "CustomResourceDefinition",
                "metadata": {"name": f"{i.plural}.kubeflow.org"},
                "spec": {"group": "kubeflow.org", "scope": "Namespaced",
                         "names": {"kind": i.kind, "plural": i.plural, "singular": i.singular},
                         "versions": [{"name": "v1", "served": True, "storage": True,
                                       "subresources": {"status": {}}}]}})

    def _on_store_event(self, resource: str, etype: str, obj) -> None:
        # a deleted job takes its rendezvous state with it: keys, generations, brokered fds all
        # live under <namespace>/<job>/<uid>/ (launcher/rendezvous.py)
        if etype == "DELETED" and resource in self._job_plurals:
            md = obj.get("metadata") or {}
            self.rendezvous.delete_prefix(
                f"{md.get('namespace', 'default')}/{md.get('name')}/{md.get('uid', '')}/")

    def start(self) -> "LocalCluster":
        if self.running:
            return self
        from .api import KINDS
        self._job_plurals = {m.INFO.plural for m in KINDS.values()}
        self.store.add_listener(self._on_store_event)
        self.rendezvous.start()
        if self.api is not None:
            self.api.start()
            self.kubelet.opt.extra_env.setdefault("TFOP_APISERVER", self.api.url)
        self.kubelet.start()
        self.manager.start()
        self.running = True
        return self

    def stop(self) -> None:
        if not self.running:
            return
        self.manager.stop()
        self.kubelet.stop()
        if self.api is not None:
            self.api.stop()
        self.store.remove_listener(self._on_store_event)
        self.rendezvous.stop()
        self.running = False

    def __enter__(self):
        return self.start()

    def __exit__(self, *exc):
        self.stop()


def default_cluster() -> LocalCluster:
    """Process-wide cluster used by SDK clients created without an explicit target."""
    global _default
    with _default_lock:
        if _default is None or not _default.running:
            _default = LocalCluster()
        return _default


def reset_default_cluster() -> None:
    global _default
    with _default_lock:
        if _default is not None:
            _default.stop()
        _default = None
