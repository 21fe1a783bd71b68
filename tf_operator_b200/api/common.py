"""Shared API vocabulary of the four job kinds (group kubeflow.org/v1).

Objects are plain JSON-shaped dicts with the reference's camelCase keys, so a
YAML written for the reference loads unchanged and what we write back is
wire-compatible. This module holds the constants and small helpers of the
absent ``kubeflow/common`` v1 API package as the reference uses them
(schema: manifests/base/kubeflow.org_tfjobs.yaml:36-108
and :6816-6896).
"""
from __future__ import annotations

import copy
import pickle as _pickle
import datetime as _dt
from dataclasses import dataclass
from typing import Any, Dict, Iterable, List, Optional, Tuple

GROUP = "kubeflow.org"
VERSION = "v1"
API_VERSION = f"{GROUP}/{VERSION}"

# labels (SDK constants: sdk/python/kubeflow/tfjob/constants/constants.py:29-33)
LABEL_GROUP_NAME = "group-name"
LABEL_JOB_NAME = "job-name"
LABEL_REPLICA_TYPE = "replica-type"
LABEL_REPLICA_INDEX = "replica-index"
LABEL_JOB_ROLE = "job-role"

# gang scheduling (pkg/controller.v1/tensorflow/pod.go:42,54,235-236)
GANG_SCHEDULER_NAME = "volcano"
ANNOTATION_POD_GROUP = "scheduling.k8s.io/group-name"
ANNOTATION_VOLCANO_TASK_SPEC = "volcano.sh/task-spec"

# restart policies
RESTART_ALWAYS, RESTART_ON_FAILURE, RESTART_NEVER, RESTART_EXIT_CODE = (
    "Always", "OnFailure", "Never", "ExitCode")
RESTART_POLICIES = (RESTART_ALWAYS, RESTART_ON_FAILURE, RESTART_NEVER, RESTART_EXIT_CODE)

# clean pod policies
CLEAN_ALL, CLEAN_RUNNING, CLEAN_NONE = "All", "Running", "None"

# job condition types
JOB_CREATED, JOB_RUNNING, JOB_RESTARTING, JOB_SUCCEEDED, JOB_FAILED = (
    "Created", "Running", "Restarting", "Succeeded", "Failed")

# pod phases
POD_PENDING, POD_RUNNING, POD_SUCCEEDED, POD_FAILED, POD_UNKNOWN = (
    "Pending", "Running", "Succeeded", "Failed", "Unknown")

# common condition reasons (commonutil.Job*Reason)
REASON_JOB_RUNNING = "JobRunning"
REASON_JOB_SUCCEEDED = "JobSucceeded"
REASON_JOB_FAILED = "JobFailed"
REASON_JOB_RESTARTING = "JobRestarting"

ENV_KUBEFLOW_NAMESPACE = "KUBEFLOW_NAMESPACE"      # pkg/common/constants.go:5
ENV_CUSTOM_CLUSTER_DOMAIN = "CUSTOM_CLUSTER_DOMAIN"  # tensorflow.go:29-33


class ValidationError(ValueError):
    """Spec failed validation (pkg/apis/*/validation/validation.go)."""


@dataclass(frozen=True)
class KindInfo:
    kind: str
    plural: str
    singular: str
    specs_key: str                 # tfReplicaSpecs / pytorchReplicaSpecs / ...
    container: str                 # default container name
    port_name: str
    port: int
    restart_policy: str            # default replica restartPolicy
    clean_pod_policy: str          # default runPolicy.cleanPodPolicy
    replica_types: Tuple[str, ...]  # canonical spellings
    normalised_types: Tuple[str, ...]  # spellings fixed up case-insensitively by defaulting
    controller_name: str


def now_rfc3339() -> str:
    return _dt.datetime.now(_dt.timezone.utc).strftime("%Y-%m-%dT%H:%M:%SZ")


def parse_time(ts: str) -> float:
    """RFC3339 (second or sub-second precision) -> unix seconds."""
    ts = ts.replace("Z", "+00:00")
    return _dt.datetime.fromisoformat(ts).timestamp()


def deepcopy(obj):
    """Deep copy of a JSON-shaped object. API objects are plain dict / list / str / int / bool
    trees, for which a pickle round trip is ~2.5x faster than copy.deepcopy (29 vs 72 us for a
    pod) - and the API server copies on every read and every watch event, under the GIL, so
    this is the control plane's hottest function. Anything pickle cannot take falls back."""
    try:
        return _pickle.loads(_pickle.dumps(obj, _pickle.HIGHEST_PROTOCOL))
    except Exception:  # noqa: BLE001 - exotic value: keep the general path
        return copy.deepcopy(obj)


def meta(obj: Dict[str, Any]) -> Dict[str, Any]:
    return obj.setdefault("metadata", {})


def name_of(obj) -> str:
    return obj.get("metadata", {}).get("name", "")


def namespace_of(obj) -> str:
    return obj.get("metadata", {}).get("namespace", "") or "default"


def key_of(obj) -> str:
    """``<namespace>/<name>`` (common.KeyFunc)."""
    return f"{namespace_of(obj)}/{name_of(obj)}"


def gen_general_name(job_name: str, rtype: str, index) -> str:
    """``<job>-<lower(type)>-<index>`` with '/' -> '-' (common.GenGeneralName;
    local copy at pkg/controller.v1/pytorch/pytorch.go:78-81)."""
    return f"{job_name}-{rtype.lower()}-{index}".replace("/", "-")


def gen_labels(job_name: str) -> Dict[str, str]:
    """Labels every owned pod/service carries (common.GenLabels)."""
    return {LABEL_GROUP_NAME: GROUP, LABEL_JOB_NAME: job_name.replace("/", "-")}


def gen_owner_reference(job: Dict[str, Any]) -> Dict[str, Any]:
    """controller_test.go:271-287."""
    return {"apiVersion": job.get("apiVersion", API_VERSION), "kind": job["kind"],
            "name": name_of(job), "uid": job.get("metadata", {}).get("uid", ""),
            "controller": True, "blockOwnerDeletion": True}


def get_controller_of(obj) -> Optional[Dict[str, Any]]:
    for ref in obj.get("metadata", {}).get("ownerReferences", []) or []:
        if ref.get("controller"):
            return ref
    return None


def replica_specs(job: Dict[str, Any], info: KindInfo) -> Dict[str, Dict[str, Any]]:
    return (job.get("spec") or {}).get(info.specs_key) or {}


def run_policy(job: Dict[str, Any]) -> Dict[str, Any]:
    return (job.get("spec") or {}).get("runPolicy") or {}


def containers_of(replica_spec) -> List[Dict[str, Any]]:
    return (((replica_spec or {}).get("template") or {}).get("spec") or {}).get("containers") or []


def total_replicas(specs: Dict[str, Dict[str, Any]]) -> int:
    return sum(int(s.get("replicas", 1) if s.get("replicas") is not None else 1)
               for s in specs.values() if s)


# ----------------------------------------------------------------------------
# flat (pre-RunPolicy) spelling -> runPolicy   
# ----------------------------------------------------------------------------
_FLAT_RUN_POLICY_KEYS = ("cleanPodPolicy", "ttlSecondsAfterFinished", "activeDeadlineSeconds",
                         "backoffLimit", "schedulingPolicy")


def lift_flat_run_policy(job: Dict[str, Any]) -> None:
    """Accept the old flat spelling some reference examples still use
    (examples/v1/mnist_with_summaries/tf_job_mnist.yaml:7) and always emit runPolicy."""
    spec = job.setdefault("spec", {})
    rp = spec.get("runPolicy")
    for k in _FLAT_RUN_POLICY_KEYS:
        if k in spec:
            val = spec.pop(k)
            if rp is None:
                rp = spec.setdefault("runPolicy", {})
            rp.setdefault(k, val)


# ----------------------------------------------------------------------------
# defaults shared by all four kinds
# ----------------------------------------------------------------------------
def _normalise_type_names(specs: Dict[str, Any], canon: Iterable[str]) -> None:
    """setTypeNamesToCamelCase (tensorflow/v1/defaults.go:72-91): ps -> PS, WORKER -> Worker."""
    for typ in canon:
        for t in list(specs.keys()):
            if t.lower() == typ.lower() and t != typ:
                specs[typ] = specs.pop(t)
                break


def _set_default_port(pod_spec: Dict[str, Any], info: KindInfo) -> None:
    """setDefaultPort (tensorflow/v1/defaults.go:38-60): append the default port to the
    default container (else container 0) unless a port with the default NAME exists."""
    containers = pod_spec.get("containers") or []
    if not containers:
        return
    index = 0
    for i, c in enumerate(containers):
        if c.get("name") == info.container:
            index = i
            break
    ports = containers[index].get("ports") or []
    if not any(p.get("name") == info.port_name for p in ports):
        ports = list(ports) + [{"name": info.port_name, "containerPort": info.port}]
        containers[index]["ports"] = ports


def _set_default_replicas(rspec: Dict[str, Any], info: KindInfo) -> None:
    if rspec.get("replicas") is None:
        rspec["replicas"] = 1
    if not rspec.get("restartPolicy"):
        rspec["restartPolicy"] = info.restart_policy


def set_common_defaults(job: Dict[str, Any], info: KindInfo, port_types: Optional[Iterable[str]] = None) -> None:
    """The part of SetDefaults_* every kind shares."""
    lift_flat_run_policy(job)
    spec = job.setdefault("spec", {})
    rp = spec.setdefault("runPolicy", {})
    if rp.get("cleanPodPolicy") is None:
        rp["cleanPodPolicy"] = info.clean_pod_policy
    specs = spec.get(info.specs_key)
    if not specs:
        return
    _normalise_type_names(specs, info.normalised_types)
    for rtype, rspec in specs.items():
        if rspec is None:
            continue
        _set_default_replicas(rspec, info)
        if port_types is None or rtype in port_types:
            pod_spec = rspec.setdefault("template", {}).setdefault("spec", {})
            _set_default_port(pod_spec, info)


def get_port(job: Dict[str, Any], info: KindInfo, rtype: str) -> Optional[int]:
    """Port named info.port_name on the default container of replica type ``rtype``."""
    rspec = replica_specs(job, info).get(rtype)
    for c in containers_of(rspec):
        if c.get("name") == info.container:
            for p in c.get("ports") or []:
                if p.get("name") == info.port_name:
                    return int(p.get("containerPort"))
    return None


# ----------------------------------------------------------------------------
# status / conditions   (commonutil.UpdateJobConditions; status_test.go:585-592)
# ----------------------------------------------------------------------------
def new_status() -> Dict[str, Any]:
    return {"conditions": [], "replicaStatuses": {}}


def status_of(job: Dict[str, Any]) -> Dict[str, Any]:
    st = job.setdefault("status", {})
    st.setdefault("conditions", [])
    st.setdefault("replicaStatuses", {})
    if st["conditions"] is None:
        st["conditions"] = []
    if st["replicaStatuses"] is None:
        st["replicaStatuses"] = {}
    return st


def get_condition(status: Dict[str, Any], ctype: str) -> Optional[Dict[str, Any]]:
    for c in status.get("conditions") or []:
        if c.get("type") == ctype:
            return c
    return None


def has_condition(status: Dict[str, Any], ctype: str) -> bool:
    c = get_condition(status, ctype)
    return c is not None and c.get("status") == "True"


def is_succeeded(status) -> bool:
    return has_condition(status, JOB_SUCCEEDED)


def is_failed(status) -> bool:
    return has_condition(status, JOB_FAILED)


def is_finished(status) -> bool:
    return is_succeeded(status) or is_failed(status)


def _filter_out(conditions: List[Dict[str, Any]], ctype: str) -> List[Dict[str, Any]]:
    out = []
    for c in conditions:
        if ctype == JOB_RESTARTING and c["type"] == JOB_RUNNING:
            continue
        if ctype == JOB_RUNNING and c["type"] == JOB_RESTARTING:
            continue
        if c["type"] == ctype:
            continue
        # a terminal condition turns an existing Running (and Failed) to False
        if ctype in (JOB_FAILED, JOB_SUCCEEDED) and c["type"] in (JOB_RUNNING, JOB_FAILED):
            c = dict(c)
            c["status"] = "False"
        out.append(c)
    return out


def update_job_conditions(status: Dict[str, Any], ctype: str, reason: str, message: str,
                          now: Optional[str] = None) -> None:
    """Append/refresh a condition with the list algebra of kubeflow/common:
    no-op once Failed; no-op for an identical (type,status,reason); the new
    condition goes LAST; Running and Restarting replace each other; a terminal
    condition flips Running to "False"; lastTransitionTime survives an
    unchanged status."""
    now = now or now_rfc3339()
    conds = status.setdefault("conditions", [])
    if conds is None:
        conds = status["conditions"] = []
    if is_failed(status):
        return
    new = {"type": ctype, "status": "True", "reason": reason, "message": message,
           "lastUpdateTime": now, "lastTransitionTime": now}
    cur = get_condition(status, ctype)
    if cur is not None and cur.get("status") == new["status"] and cur.get("reason") == reason:
        return
    if cur is not None and cur.get("status") == new["status"]:
        new["lastTransitionTime"] = cur.get("lastTransitionTime", now)
    status["conditions"] = _filter_out(conds, ctype) + [new]


def init_replica_status(status: Dict[str, Any], rtype: str) -> None:
    status.setdefault("replicaStatuses", {})[rtype] = {}


def count_pod_phase(status: Dict[str, Any], rtype: str, pod: Dict[str, Any]) -> None:
    """updateJobReplicaStatuses (status.go:253-262). Zero counters are omitted on the wire
    (Go `omitempty`), so they are only created when incremented."""
    rs = status.setdefault("replicaStatuses", {}).setdefault(rtype, {})
    phase = (pod.get("status") or {}).get("phase")
    if phase == POD_RUNNING:
        rs["active"] = rs.get("active", 0) + 1
    elif phase == POD_SUCCEEDED:
        rs["succeeded"] = rs.get("succeeded", 0) + 1
    elif phase == POD_FAILED:
        rs["failed"] = rs.get("failed", 0) + 1


# exit codes (train_util.IsRetryableExitCode; docs/design/tf_job_design_doc.md:84)
_PERMANENT = {1, 2, 126, 127, 128, 139}
_RETRYABLE = {130, 137, 143, 138}


def is_retryable_exit_code(code: int) -> bool:
    if code in _PERMANENT:
        return False
    return code in _RETRYABLE
