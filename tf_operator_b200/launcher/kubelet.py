"""Single-box launcher: the "kubelet + scheduler" that turns Pod objects into OS processes,
each replica pinned to one B200.

What the reference delegates to Kubernetes (pod scheduling, container start, restartPolicy
Always/OnFailure/Never handled by kubelet in place, phases, container exit codes, logs,
Volcano's all-or-nothing PodGroup admission — SURVEY.md 5.3, Appendix A) happens here:

  Pod ADDED  -> (gang pods wait for their PodGroup: minMember pods present AND that many GPUs
                free, then all are admitted at once; others are admitted immediately)
             -> GPU picked by NVLink topology (launcher/topology.py), env completed
                (CUDA_VISIBLE_DEVICES, TFOP_* rendezvous / namespace / pod identity)
             -> one process per container, status.phase Running, containerStatuses filled
  exit       -> restartPolicy Always/OnFailure: restarted in place with back-off
                (restartCount grows, pod object survives); otherwise phase Succeeded/Failed
                with state.terminated.exitCode — which is what the reconcilers read
  Pod DELETED-> process group SIGTERM, then SIGKILL after the grace period

"image" is recorded but not pulled: containers are `command` + `args` run on the host.
"""
from __future__ import annotations

import logging
import os
import shlex
import signal
import subprocess
import sys
import tempfile
import threading
import time
from dataclasses import dataclass, field
from pathlib import Path
from typing import Any, Dict, List, Optional

from ..api import common as c
from ..store.memory import ApiError, ObjectStore
from . import netmap
from .topology import Topology, probe, synthetic

log = logging.getLogger("tf_operator_b200.launcher")

GPU_RESOURCE = "nvidia.com/gpu"
ANNOTATION_GPUS = "tfop.b200/gpus"


@dataclass
class KubeletOptions:
    log_dir: str = ""
    grace_period: float = 2.0
    restart_backoff: float = 0.2            # first in-place restart delay, doubles up to 10 s
    gpus_per_replica_default: int = 0       # GPUs for pods that do not ask for nvidia.com/gpu
    topology: Optional[Topology] = None
    rendezvous_addr: str = ""               # exported as TFOP_RENDEZVOUS
    rendezvous_uds: str = ""                # descriptor broker socket, TFOP_RENDEZVOUS_UDS
    cluster_id: str = ""                    # mixed into every pod's loopback address (netmap)
    extra_env: Dict[str, str] = field(default_factory=dict)
    python: str = sys.executable
    cwd: str = ""
    # local "image registry": image name -> directory standing in for the image's root file
    # system. Absolute paths in command / args / `--flag=/path` whose first component exists
    # under that directory are rewritten into it, so a manifest written for a container image
    # (`python /var/tf_mnist/mnist_with_summaries.py --log_dir=/train/logs`) runs unchanged.
    image_roots: Dict[str, str] = field(default_factory=dict)


class _Container:
    def __init__(self, name: str):
        self.name = name
        self.proc: Optional[subprocess.Popen] = None
        self.restart_count = 0
        self.started_at = ""
        self.finished_at = ""
        self.exit_code: Optional[int] = None
        self.log_path: Optional[Path] = None


class _PodRuntime:
    def __init__(self, pod: Dict[str, Any]):
        self.ns, self.name = c.namespace_of(pod), c.name_of(pod)
        self.uid = pod["metadata"].get("uid")
        self.containers: Dict[str, _Container] = {}
        self.gpus: List[int] = []
        self.deleted = False
        self.lock = threading.Lock()


class _HostGpuLocks:
    """Host-wide claim on physical GPUs: one lock file per GPU (by UUID) under the system temp
    directory, held with flock for as long as a pod of THIS launcher runs on it. Two operators on
    one box (two LocalClusters, two test sessions) each keep their own free list; the lock is
    what stops them from handing the same B200 to two replicas. Synthetic topologies (tests
    without GPUs) take no locks."""

    def __init__(self, topology: Topology):
        self._enabled = topology.source not in ("none", "synthetic") and topology.count > 0
        self._names = {g.index: (g.uuid or f"index-{g.index}").replace("/", "_") for g in topology.gpus}
        self._held: Dict[int, Any] = {}
        self._dir = Path(tempfile.gettempdir()) / "tfop-b200-gpu-locks"

    def acquire(self, gpu: int) -> bool:
        if not self._enabled or gpu in self._held:
            return True
        import fcntl
        self._dir.mkdir(parents=True, exist_ok=True)
        f = open(self._dir / f"{self._names.get(gpu, gpu)}.lock", "a+")
        try:
            fcntl.flock(f, fcntl.LOCK_EX | fcntl.LOCK_NB)
        except OSError:
            f.close()
            return False
        f.seek(0)
        f.truncate()
        f.write(f"{os.getpid()}\n")
        f.flush()
        self._held[gpu] = f
        return True

    def release(self, gpu: int) -> None:
        f = self._held.pop(gpu, None)
        if f is not None:
            f.close()                      # closing the descriptor drops the flock


class LocalKubelet:
    def __init__(self, store: ObjectStore, options: Optional[KubeletOptions] = None):
        self.store = store
        self.opt = options or KubeletOptions()
        self.topology = self.opt.topology if self.opt.topology is not None else probe()
        self._free_gpus = [g.index for g in self.topology.gpus]
        self._gpu_locks = _HostGpuLocks(self.topology)
        self._pods: Dict[tuple, _PodRuntime] = {}
        self._pending_gang: Dict[tuple, Dict[str, Any]] = {}
        self._unschedulable: Dict[tuple, tuple] = {}    # (ns, name) -> (pod, group): retried on release
        self._lock = threading.RLock()
        self._stop = threading.Event()
        self._log_dir = Path(self.opt.log_dir or tempfile.mkdtemp(prefix="tfop-logs-"))
        self._log_dir.mkdir(parents=True, exist_ok=True)
        self.started = False

    # ------------------------------------------------------------------ lifecycle
    def start(self) -> "LocalKubelet":
        self.store.add_listener(self._on_event)
        for pod in self.store.list("pods"):
            self._on_event("pods", "ADDED", pod)
        self.started = True
        return self

    def stop(self) -> None:
        self._stop.set()
        self.store.remove_listener(self._on_event)
        with self._lock:
            pods = list(self._pods.values())
        for rt in pods:
            self._kill(rt)
        self.started = False

    # ------------------------------------------------------------------ events
    def _on_event(self, resource: str, etype: str, obj) -> None:
        if self._stop.is_set():
            return
        if resource == "podgroups" and etype in ("ADDED", "MODIFIED"):
            self._try_admit_gangs()
            return
        if resource != "pods":
            return
        key = (c.namespace_of(obj), c.name_of(obj))
        if etype == "ADDED":
            if (obj.get("status") or {}).get("phase", c.POD_PENDING) != c.POD_PENDING:
                return
            if self._is_gang_pod(obj):
                with self._lock:
                    self._pending_gang[key] = obj
                self._try_admit_gangs()
            else:
                self._admit([obj])
        elif etype == "DELETED":
            with self._lock:
                self._pending_gang.pop(key, None)
                self._unschedulable.pop(key, None)
                rt = self._pods.pop(key, None)
            if rt is not None and rt.uid == obj["metadata"].get("uid"):
                threading.Thread(target=self._kill, args=(rt,), daemon=True).start()
            self._try_admit_gangs()

    # ------------------------------------------------------------------ scheduling
    @staticmethod
    def _is_gang_pod(pod) -> bool:
        ann = pod.get("metadata", {}).get("annotations") or {}
        return bool(ann.get(c.ANNOTATION_POD_GROUP)) and \
            (pod.get("spec") or {}).get("schedulerName") == c.GANG_SCHEDULER_NAME

    def _gpus_wanted(self, pod) -> int:
        n = 0
        for ctr in (pod.get("spec") or {}).get("containers") or []:
            lim = ((ctr.get("resources") or {}).get("limits") or {}).get(GPU_RESOURCE)
            if lim is not None:
                n += int(lim)
        return n if n > 0 else self.opt.gpus_per_replica_default

    def _try_admit_gangs(self) -> None:
        with self._lock:
            groups: Dict[tuple, List[Dict[str, Any]]] = {}
            for (ns, _), pod in self._pending_gang.items():
                g = pod["metadata"]["annotations"][c.ANNOTATION_POD_GROUP]
                groups.setdefault((ns, g), []).append(pod)
            for (ns, g), pods in groups.items():
                try:
                    pg = self.store.get("podgroups", ns, g)
                except ApiError:
                    continue
                min_member = int((pg.get("spec") or {}).get("minMember", len(pods)))
                running = sum(1 for (pns, _), rt in self._pods.items() if pns == ns and
                              getattr(rt, "group", None) == g)
                if len(pods) + running < min_member:
                    continue
                need = sum(self._gpus_wanted(p) for p in pods)
                if need > len(self._free_gpus):
                    self._set_podgroup_phase(pg, "Pending",
                                             f"{need} GPUs wanted, {len(self._free_gpus)} free")
                    continue
                for p in pods:
                    self._pending_gang.pop((ns, c.name_of(p)), None)
                self._set_podgroup_phase(pg, "Running", "")
                self._admit(pods, group=g)

    def _set_podgroup_phase(self, pg, phase: str, message: str) -> None:
        st = pg.setdefault("status", {})
        if st.get("phase") == phase:
            return
        st["phase"] = phase
        if message:
            st["message"] = message
        try:
            self.store.update_status("podgroups", pg)
        except ApiError:
            pass

    def _admit(self, pods: List[Dict[str, Any]], group: Optional[str] = None) -> None:
        with self._lock:
            for pod in pods:
                want = self._gpus_wanted(pod)
                gpus = self._claim_gpus(want) if want else []
                if gpus is None:
                    # stays Pending with PodScheduled=False, like a pod the scheduler cannot
                    # place; tried again whenever a pod gives its GPUs back
                    self._unschedulable[(c.namespace_of(pod), c.name_of(pod))] = (pod, group)
                    self._fail_unschedulable(pod, f"Insufficient {GPU_RESOURCE}: want {want}, "
                                                  f"free {len(self._free_gpus)}")
                    continue
                self._unschedulable.pop((c.namespace_of(pod), c.name_of(pod)), None)
                for g in gpus:
                    self._free_gpus.remove(g)
                rt = _PodRuntime(pod)
                rt.gpus = gpus
                rt.group = group
                self._pods[(rt.ns, rt.name)] = rt
                threading.Thread(target=self._run_pod, args=(rt, pod), daemon=True,
                                 name=f"pod-{rt.name}").start()

    def _claim_gpus(self, want: int) -> Optional[List[int]]:
        """Pick `want` free GPUs by topology and take their host-wide locks; a GPU another
        launcher on this host holds is skipped. All or nothing."""
        candidates = list(self._free_gpus)
        while True:
            gpus = self.topology.pick(want, candidates)
            if gpus is None:
                return None
            got = []
            for g in gpus:
                if self._gpu_locks.acquire(g):
                    got.append(g)
                else:
                    candidates.remove(g)
            if len(got) == len(gpus):
                return gpus
            for g in got:
                self._gpu_locks.release(g)

    def _fail_unschedulable(self, pod, message: str) -> None:
        try:
            pod = self.store.get("pods", c.namespace_of(pod), c.name_of(pod))
        except ApiError:
            return
        conds = (pod.get("status") or {}).get("conditions") or []
        if any(x.get("reason") == "Unschedulable" and x.get("message") == message for x in conds):
            return
        st = pod.setdefault("status", {})
        st["phase"] = c.POD_PENDING
        st["conditions"] = [{"type": "PodScheduled", "status": "False", "reason": "Unschedulable",
                             "message": message}]
        try:
            self.store.update_status("pods", pod)
        except ApiError:
            pass

    # ------------------------------------------------------------------ running
    def _container_cmd(self, ctr) -> List[str]:
        cmd = list(ctr.get("command") or []) + list(ctr.get("args") or [])
        if not cmd:
            raise ValueError(f"container {ctr.get('name')} has neither command nor args "
                             f"(images are not pulled by the local launcher)")
        if cmd[0] in ("python", "python3"):
            cmd[0] = self.opt.python
        root = self.opt.image_roots.get(ctr.get("image") or "")
        if root:
            cmd = [cmd[0]] + [self._into_image_root(str(a), root) for a in cmd[1:]]
        return [str(x) for x in cmd]

    @staticmethod
    def _into_image_root(arg: str, root: str) -> str:
        head, sep, tail = arg.partition("=")
        path, prefix = (tail, head + sep) if (sep and tail.startswith("/")) else (arg, "")
        if not path.startswith("/"):
            return arg
        first = path.lstrip("/").split("/", 1)[0]
        if first and (Path(root) / first).exists():
            return prefix + str(Path(root) / path.lstrip("/"))
        return arg

    def _env_for(self, rt: _PodRuntime, pod, ctr) -> Dict[str, str]:
        env = dict(os.environ)
        env.update(self.opt.extra_env)
        for e in ctr.get("env") or []:
            if "value" in e:
                env[e["name"]] = str(e["value"])
        labels = pod["metadata"].get("labels") or {}
        env.update({
            netmap.ENV_NAMESPACE: rt.ns, "TFOP_POD_NAME": rt.name,
            "TFOP_POD_IP": netmap.pod_ip(rt.ns, rt.name, self.opt.cluster_id),
            netmap.ENV_CLUSTER_ID: self.opt.cluster_id,
            "TFOP_JOB_NAME": labels.get(c.LABEL_JOB_NAME, ""),
            "TFOP_REPLICA_TYPE": labels.get(c.LABEL_REPLICA_TYPE, ""),
            "TFOP_REPLICA_INDEX": labels.get(c.LABEL_REPLICA_INDEX, ""),
            "TFOP_GPU_COUNT": str(len(rt.gpus)),
            # identity of this incarnation: the owning job's uid scopes every rendezvous key (a
            # re-created job of the same name starts clean); pod uid + restart count tell a
            # restarted replica apart from the process it replaces
            "TFOP_JOB_UID": next((o.get("uid", "") for o in
                                  pod["metadata"].get("ownerReferences") or []
                                  if o.get("controller")), ""),
            "TFOP_POD_UID": str(rt.uid or ""),
            "TFOP_INCARNATION": str(rt.containers[ctr["name"]].restart_count
                                    if ctr["name"] in rt.containers else 0),
        })
        if self.opt.rendezvous_addr:
            env["TFOP_RENDEZVOUS"] = self.opt.rendezvous_addr
        if self.opt.rendezvous_uds:
            env["TFOP_RENDEZVOUS_UDS"] = self.opt.rendezvous_uds
        if rt.gpus:
            env["CUDA_VISIBLE_DEVICES"] = ",".join(str(g) for g in rt.gpus)
        elif self.topology.count and self._gpus_wanted(pod) == 0 and \
                not self.opt.gpus_per_replica_default:
            pass  # CPU pod on a GPU box: leave visibility as inherited
        repo_root = str(Path(__file__).resolve().parents[2])
        env["PYTHONPATH"] = repo_root + (os.pathsep + env["PYTHONPATH"] if env.get("PYTHONPATH") else "")
        return env

    def _start_container(self, rt: _PodRuntime, pod, ctr) -> _Container:
        cs = rt.containers.get(ctr["name"]) or _Container(ctr["name"])
        rt.containers[ctr["name"]] = cs
        cs.log_path = self._log_dir / rt.ns / rt.name / f"{ctr['name']}.log"
        cs.log_path.parent.mkdir(parents=True, exist_ok=True)
        logf = open(cs.log_path, "ab", buffering=0)
        cwd = ctr.get("workingDir") or self.opt.cwd or None
        cs.proc = subprocess.Popen(self._container_cmd(ctr), env=self._env_for(rt, pod, ctr),
                                   stdout=logf, stderr=subprocess.STDOUT, cwd=cwd,
                                   start_new_session=True)
        logf.close()
        cs.started_at = c.now_rfc3339()
        cs.exit_code = None
        return cs

    def _publish(self, rt: _PodRuntime, phase: str, reason: str = "", message: str = "") -> None:
        statuses = []
        for cs in rt.containers.values():
            if cs.exit_code is None and cs.proc is not None:
                state = {"running": {"startedAt": cs.started_at}}
            elif cs.exit_code is not None:
                state = {"terminated": {"exitCode": cs.exit_code, "startedAt": cs.started_at,
                                        "finishedAt": cs.finished_at,
                                        "reason": "Completed" if cs.exit_code == 0 else "Error"}}
            else:
                state = {"waiting": {"reason": reason or "ContainerCreating"}}
            statuses.append({"name": cs.name, "state": state, "restartCount": cs.restart_count,
                             "ready": "running" in state})
        try:
            pod = self.store.get("pods", rt.ns, rt.name)
        except ApiError:
            return
        if pod["metadata"].get("uid") != rt.uid:
            return
        st = pod.setdefault("status", {})
        st.update({"phase": phase, "containerStatuses": statuses,
                   "podIP": netmap.pod_ip(rt.ns, rt.name, self.opt.cluster_id),
                   "hostIP": "127.0.0.1"})
        st.setdefault("startTime", c.now_rfc3339())
        if reason:
            st["reason"] = reason
        if message:
            st["message"] = message
        if rt.gpus:
            pod["metadata"].setdefault("annotations", {})
        try:
            self.store.update_status("pods", pod)
        except ApiError:
            pass

    def _run_pod(self, rt: _PodRuntime, pod) -> None:
        spec = pod.get("spec") or {}
        policy = spec.get("restartPolicy") or "Always"
        containers = spec.get("containers") or []
        try:
            with rt.lock:
                if rt.deleted:
                    return
                for ctr in containers:
                    self._start_container(rt, pod, ctr)
        except Exception as e:  # noqa: BLE001
            log.error("failed to start pod %s/%s: %s", rt.ns, rt.name, e)
            for ctr in containers:
                cs = rt.containers.setdefault(ctr["name"], _Container(ctr["name"]))
                if cs.proc is None:
                    cs.exit_code, cs.finished_at = 128, c.now_rfc3339()
            self._publish(rt, c.POD_FAILED, "StartError", str(e))
            self._release(rt)
            return
        self._publish(rt, c.POD_RUNNING)
        backoff = self.opt.restart_backoff
        by_name = {ctr["name"]: ctr for ctr in containers}
        while not rt.deleted and not self._stop.is_set():
            alive = False
            changed = False
            for cs in rt.containers.values():
                if cs.proc is None or cs.exit_code is not None:
                    continue
                rc = cs.proc.poll()
                if rc is None:
                    alive = True
                    continue
                cs.exit_code = rc if rc >= 0 else 128 - rc   # killed by signal N -> 128+N
                cs.finished_at = c.now_rfc3339()
                changed = True
                restart = policy == "Always" or (policy == "OnFailure" and cs.exit_code != 0)
                if restart and not rt.deleted:
                    self._publish(rt, c.POD_RUNNING, "CrashLoopBackOff" if cs.exit_code else "")
                    if self._stop.wait(backoff):
                        return
                    backoff = min(backoff * 2, 10.0)
                    with rt.lock:
                        if rt.deleted:
                            return
                        cs.restart_count += 1
                        self._start_container(rt, pod, by_name[cs.name])
                    alive = True
            if changed and alive:
                self._publish(rt, c.POD_RUNNING)
            if not alive:
                break
            time.sleep(0.02)
        if rt.deleted:
            return
        codes = [cs.exit_code or 0 for cs in rt.containers.values()]
        phase = c.POD_SUCCEEDED if all(x == 0 for x in codes) else c.POD_FAILED
        self._publish(rt, phase)
        self._release(rt)

    def _release(self, rt: _PodRuntime) -> None:
        with self._lock:
            for g in rt.gpus:
                self._gpu_locks.release(g)
                if g not in self._free_gpus:
                    self._free_gpus.append(g)
            rt.gpus = []
            for key in list(self._unschedulable):
                pod, group = self._unschedulable.pop(key)
                self._admit([pod], group=group)      # goes back on the list if it still does not fit
        self._try_admit_gangs()

    def _kill(self, rt: _PodRuntime) -> None:
        with rt.lock:
            rt.deleted = True
            procs = [cs.proc for cs in rt.containers.values() if cs.proc is not None]
        for p in procs:
            if p.poll() is None:
                try:
                    os.killpg(p.pid, signal.SIGTERM)
                except (ProcessLookupError, PermissionError):
                    pass
        end = time.time() + self.opt.grace_period
        for p in procs:
            try:
                p.wait(timeout=max(0.0, end - time.time()))
            except subprocess.TimeoutExpired:
                try:
                    os.killpg(p.pid, signal.SIGKILL)
                except (ProcessLookupError, PermissionError):
                    pass
                p.wait()
        self._release(rt)

    # ------------------------------------------------------------------ logs / introspection
    def log_path(self, namespace: str, pod: str, container: Optional[str] = None) -> Optional[Path]:
        d = self._log_dir / namespace / pod
        if not d.exists():
            return None
        if container:
            p = d / f"{container}.log"
            return p if p.exists() else None
        files = sorted(d.glob("*.log"))
        return files[0] if files else None

    def read_log(self, namespace: str, pod: str, container: Optional[str] = None) -> str:
        p = self.log_path(namespace, pod, container)
        return p.read_text(errors="replace") if p else ""

    def follow_log(self, namespace: str, pod: str, container: Optional[str] = None,
                   poll: float = 0.05, stop: Optional[threading.Event] = None):
        """Generator of log lines until the pod finishes (kubectl logs -f)."""
        pos = 0
        buf = ""
        while True:
            p = self.log_path(namespace, pod, container)
            done = self._pod_finished(namespace, pod)
            if p is not None:
                with open(p, "r", errors="replace") as f:
                    f.seek(pos)
                    chunk = f.read()
                    pos = f.tell()
                buf += chunk
                while "\n" in buf:
                    line, buf = buf.split("\n", 1)
                    yield line
            if done or (stop is not None and stop.is_set()):
                if buf:
                    yield buf
                return
            time.sleep(poll)

    def _pod_finished(self, namespace: str, pod: str) -> bool:
        try:
            ph = (self.store.get("pods", namespace, pod).get("status") or {}).get("phase")
        except ApiError:
            return True
        return ph in (c.POD_SUCCEEDED, c.POD_FAILED)

    def free_gpus(self) -> List[int]:
        with self._lock:
            return sorted(self._free_gpus)
