"""Replica-side view of the operator's environment contract.

The operator's whole contribution to a workload is the environment it injects
(SURVEY.md Appendix A): ``TF_CONFIG`` for TFJob (pkg/controller.v1/tensorflow/tensorflow.go:
51-132), ``MASTER_ADDR/MASTER_PORT/WORLD_SIZE/RANK`` for PyTorchJob and XGBoostJob
(pytorch.go:41-65, xgboost.go:59-94), ``MX_CONFIG`` + ``DMLC_*`` for MXJob (mxnet.go:74-118).
``ReplicaContext.from_env()`` reads whichever is present and gives one normalised view, with
service names resolved to this box's per-pod loopback addresses (launcher/netmap.py).
"""
from __future__ import annotations

import json
import os
from dataclasses import dataclass, field
from typing import Dict, List, Optional

from ..launcher import netmap


@dataclass
class ReplicaContext:
    kind: str                         # "TFJob" | "PyTorchJob" | "XGBoostJob" | "MXJob" | "local"
    task_type: str = "worker"
    task_index: int = 0
    cluster: Dict[str, List[str]] = field(default_factory=dict)   # role -> ["host:port", ...]
    rank: int = 0                     # rank among the data-parallel trainers
    world_size: int = 1
    master_addr: str = "127.0.0.1"
    master_port: int = 29500
    namespace: str = "default"
    job_name: str = ""
    raw: Dict[str, str] = field(default_factory=dict)

    # ---- derived ---------------------------------------------------------------
    @property
    def is_chief(self) -> bool:
        """chief/master replica, else worker 0 (IsMasterRole, tfjob_controller.go:586-593)."""
        if self.kind == "TFJob":
            if "chief" in self.cluster or "master" in self.cluster:
                return self.task_type in ("chief", "master")
            return self.task_type == "worker" and self.task_index == 0
        return self.rank == 0

    def resolved(self, role: str) -> List[str]:
        return [netmap.resolve_endpoint(e, self.namespace) for e in self.cluster.get(role, [])]

    def my_endpoint(self) -> Optional[str]:
        lst = self.cluster.get(self.task_type, [])
        if self.task_index < len(lst):
            return netmap.resolve_endpoint(lst[self.task_index], self.namespace)
        return None

    @property
    def trainers(self) -> List[str]:
        """Roles that take part in data-parallel training, in rank order."""
        return [r for r in ("chief", "master", "worker") if r in self.cluster]

    # ---- construction ------------------------------------------------------------
    @classmethod
    def from_env(cls, env: Optional[Dict[str, str]] = None) -> "ReplicaContext":
        env = dict(os.environ if env is None else env)
        ns = env.get(netmap.ENV_NAMESPACE, "default")
        job = env.get("TFOP_JOB_NAME", "")
        if env.get("TF_CONFIG"):
            cfg = json.loads(env["TF_CONFIG"])
            task = cfg.get("task", {})
            if "sparseCluster" in cfg:
                sc = cfg["sparseCluster"]
                workers = sc.get("worker") or {}
                cluster = {"ps": list(sc.get("ps") or []),
                           "worker": [workers[k] for k in sorted(workers, key=int)]}
            else:
                cluster = {k: list(v) for k, v in (cfg.get("cluster") or {}).items()}
            ctx = cls("TFJob", task.get("type", "worker"), int(task.get("index", 0)), cluster,
                      namespace=ns, job_name=job, raw=env)
            rank, world = 0, 0
            for role in ctx.trainers:
                n = len(cluster[role])
                if role == ctx.task_type:
                    rank = world + ctx.task_index
                world += n
            ctx.rank, ctx.world_size = rank, max(world, 1)
            first = next((cluster[r][0] for r in ctx.trainers if cluster[r]), None)
            if first:
                host, _, port = first.rpartition(":")
                ctx.master_addr = netmap.resolve(host, ns)
                # the TF port (2222) belongs to the workload; torch rendezvous uses port+1
                ctx.master_port = int(port) + 1
            return ctx
        if env.get("MX_CONFIG"):
            cfg = json.loads(env["MX_CONFIG"])
            cluster = {k: [f"{e['url']}:{e['port']}" for e in v]
                       for k, v in (cfg.get("cluster") or {}).items()}
            task = cfg.get("task", {})
            ctx = cls("MXJob", task.get("type", "worker"), int(task.get("index", 0)), cluster,
                      namespace=ns, job_name=job, raw=env)
            ctx.world_size = int(env.get("DMLC_NUM_WORKER", len(cluster.get("worker", [])) or 1))
            ctx.rank = int(env.get("DMLC_WORKER_ID", ctx.task_index))
            ctx.master_addr = netmap.resolve(env.get("DMLC_PS_ROOT_URI", "127.0.0.1"), ns)
            ctx.master_port = int(env.get("DMLC_PS_ROOT_PORT", 9091))
            return ctx
        if env.get("MASTER_ADDR") and env.get("RANK") is not None:
            rank, world = int(env["RANK"]), int(env.get("WORLD_SIZE", 1))
            kind = "XGBoostJob" if env.get("TFOP_REPLICA_TYPE") and "WORKER_ADDRS" in env \
                else "PyTorchJob"
            ctx = cls(kind, env.get("TFOP_REPLICA_TYPE", "master" if rank == 0 else "worker"),
                      int(env.get("TFOP_REPLICA_INDEX", 0) or 0), {}, rank, world,
                      namespace=ns, job_name=job, raw=env)
            addr = env["MASTER_ADDR"]
            if addr == "localhost" and env.get("TFOP_POD_IP"):
                addr = env["TFOP_POD_IP"]      # the master binds its own pod address
            ctx.master_addr = netmap.resolve(addr, ns)
            ctx.master_port = int(env.get("MASTER_PORT", 23456))
            if env.get("WORKER_ADDRS"):
                ctx.cluster["worker"] = [f"{h}:{env.get('WORKER_PORT', '9999')}"
                                         for h in env["WORKER_ADDRS"].split(",") if h]
            return ctx
        return cls("local", raw=env)

    # ---- torch.distributed / symmetric memory -------------------------------------------
    def init_process_group(self, backend: str = "gloo", timeout_s: float = 120.0):
        """env:// rendezvous on the resolved master address, rank/world from the contract."""
        import datetime

        import torch.distributed as dist
        if dist.is_initialized():
            return
        dist.init_process_group(
            backend, init_method=f"tcp://{self.master_addr}:{self.master_port}",
            rank=self.rank, world_size=self.world_size,
            timeout=datetime.timedelta(seconds=timeout_s))

    def symmetric_comm(self, nbytes: int, tag: str = "comm0", nvls_bytes: int = 0):
        """Symmetric peer-memory communicator; handles are exchanged through the launcher's
        rendezvous store when there is one, else through torch.distributed. ``nvls_bytes`` > 0
        additionally binds a multicast (NVLS) buffer of that size when the box supports it
        (``comm.nvls_bytes`` is 0 otherwise); under the launcher the multicast descriptor goes
        through its broker too, so no torch.distributed group is created."""
        comm = self._symmetric_comm(nbytes, tag, nvls_bytes)
        if getattr(comm, "_rdv", None) is None:
            comm.nvls_bytes = 0
            if nvls_bytes and self.world_size > 1:
                import torch.distributed as dist
                if not dist.is_initialized():
                    self.init_process_group("gloo")
                comm.enable_nvls(nvls_bytes)
        return comm

    def rendezvous_scope(self, tag: str) -> str:
        """``<namespace>/<job>/<job uid>/<tag>``: the uid (TFOP_JOB_UID, exported by the launcher)
        keeps a re-created job of the same name away from its predecessor's keys."""
        return f"{self.namespace}/{self.job_name}/{self.raw.get('TFOP_JOB_UID', '')}/{tag}"

    def _symmetric_comm(self, nbytes: int, tag: str, nvls_bytes: int = 0):
        from .symm import SymmetricComm
        addr = self.raw.get("TFOP_RENDEZVOUS")
        if addr and self.world_size > 1:
            # IPC handles AND the multicast buffer are set up inside one rendezvous generation
            return rendezvous_comm(addr, self.raw.get("TFOP_RENDEZVOUS_UDS", ""),
                                   self.rendezvous_scope(tag), self.rank, self.world_size, nbytes,
                                   nvls_bytes=nvls_bytes)
        if self.world_size > 1:
            return SymmetricComm.from_torch_distributed(nbytes)
        return SymmetricComm.create(0, 1, nbytes, lambda b: [b])


_ABANDONED: list = []      # communicators of superseded generations (see rendezvous_comm)


def rendezvous_comm(addr: str, uds_path: str, scope: str, rank: int, world: int, nbytes: int,
                    nvls_bytes: int = 0):
    """Form a SymmetricComm through a launcher rendezvous store at ``addr`` (descriptor broker at
    ``uds_path``): join the scope's current generation (a restarted replica opens a new one, its
    peers follow once their collectives have timed out and the operator has restarted them),
    exchange the CUDA IPC handles under ``<scope>/g<generation>/``, barrier. The communicator
    keeps the client for the NVLS descriptor hand-off and for its exit barrier."""
    from ..launcher.rendezvous import RendezvousClient
    from .symm import SymmetricComm
    import time
    from ..launcher.rendezvous import Superseded
    cli = RendezvousClient(addr, uds_path=uds_path)
    deadline = time.time() + cli.timeout
    while True:
        gen = cli.join(scope, rank, world, timeout=max(1.0, deadline - time.time()))
        prefix = f"{scope}/g{gen}"
        comm = None
        try:
            # every wait names its generation: if a peer restarts while the handles are being
            # exchanged, the generation is over and this rank goes back to join the next one
            # (seen on 8 GPUs: one rank got through join just before a late restart superseded
            # the generation, waited for handles nobody would publish, and everyone timed out)
            comm = SymmetricComm.create(
                rank, world, nbytes,
                lambda b: cli.all_gather(prefix + "/ipc", rank, world, b, scope=scope, gen=gen))
            cli.barrier(prefix + "/ready", world, scope=scope, gen=gen)
            comm._rdv = (cli, prefix)
            comm._rdv_gen = (scope, gen)
            comm.nvls_bytes = 0
            if nvls_bytes:
                comm.enable_nvls(nvls_bytes)
            break
        except Superseded:
            if comm is not None:
                _ABANDONED.append(comm)      # peers may have mapped it: kept until process exit
            if time.time() >= deadline:
                raise TimeoutError("rendezvous kept being superseded while exchanging handles")
    comm.generation = gen
    comm._exit_barrier = lambda: cli.barrier(prefix + "/exit", world, timeout=60)
    return comm
