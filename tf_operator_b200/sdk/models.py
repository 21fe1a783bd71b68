"""SDK model objects: V1TFJob & friends plus the PyTorchJob / MXJob / XGBoostJob analogues and
light stand-ins for the few ``kubernetes.client`` models a job spec needs.

Same surface as the reference's swagger-generated classes (sdk/python/kubeflow/tfjob/models/
*.py: ``swagger_types`` / ``attribute_map``, ``to_dict()``, ``to_str()``, ``__eq__``), written
once as a small declarative base instead of 1.7k generated lines. ``to_wire()`` produces the
camelCase JSON the API server takes. V1TFJobSpec keeps the reference SDK's flat pre-RunPolicy
fields (models/v1_tf_job_spec.py:49-63) AND accepts ``run_policy``; the API server lifts the
flat spelling into runPolicy.
"""
from __future__ import annotations

import pprint
import re
from typing import Any, Dict


def _camel(s: str) -> str:
    parts = s.split("_")
    return parts[0] + "".join(p.title() if p != "ip" else "IP" for p in parts[1:])


def _snake(s: str) -> str:
    return re.sub(r"(?<!^)(?=[A-Z])", "_", s).lower()


def to_wire(value):
    """Model / dict / list -> plain JSON with camelCase keys, None fields dropped."""
    if isinstance(value, Model):
        out = {}
        for attr, key in value.attribute_map.items():
            v = getattr(value, attr)
            if v is not None:
                out[key] = to_wire(v)
        return out
    if isinstance(value, dict):
        return {k: to_wire(v) for k, v in value.items() if v is not None}
    if isinstance(value, (list, tuple)):
        return [to_wire(v) for v in value]
    return value


class Model:
    swagger_types: Dict[str, str] = {}
    attribute_map: Dict[str, str] = {}

    def __init__(self, **kwargs):
        for attr in self.swagger_types:
            setattr(self, attr, kwargs.pop(attr, None))
        if kwargs:
            raise TypeError(f"{type(self).__name__} got unexpected fields {sorted(kwargs)}")

    def to_dict(self) -> Dict[str, Any]:
        def conv(v):
            if isinstance(v, Model):
                return v.to_dict()
            if isinstance(v, list):
                return [conv(x) for x in v]
            if isinstance(v, dict):
                return {k: conv(x) for k, x in v.items()}
            return v
        return {attr: conv(getattr(self, attr)) for attr in self.swagger_types}

    def to_wire(self) -> Dict[str, Any]:
        return to_wire(self)

    def to_str(self) -> str:
        return pprint.pformat(self.to_dict())

    def __repr__(self) -> str:
        return self.to_str()

    def __eq__(self, other) -> bool:
        return isinstance(other, type(self)) and self.__dict__ == other.__dict__

    def __ne__(self, other) -> bool:
        return not self == other


def _model(name: str, fields: Dict[str, str], overrides: Dict[str, str] | None = None, doc: str = ""):
    amap = {f: (overrides or {}).get(f, _camel(f)) for f in fields}
    return type(name, (Model,), {"swagger_types": dict(fields), "attribute_map": amap,
                                 "__doc__": doc})


# ---- kubernetes.client stand-ins -----------------------------------------------------
V1ObjectMeta = _model("V1ObjectMeta", {
    "name": "str", "namespace": "str", "labels": "dict(str, str)", "annotations": "dict(str, str)",
    "generate_name": "str", "uid": "str", "resource_version": "str",
    "creation_timestamp": "datetime", "owner_references": "list[object]"})
V1ContainerPort = _model("V1ContainerPort", {"name": "str", "container_port": "int",
                                             "protocol": "str"})
V1EnvVar = _model("V1EnvVar", {"name": "str", "value": "str"})
V1ResourceRequirements = _model("V1ResourceRequirements", {"limits": "dict(str, str)",
                                                           "requests": "dict(str, str)"})
V1Container = _model("V1Container", {
    "name": "str", "image": "str", "command": "list[str]", "args": "list[str]",
    "env": "list[V1EnvVar]", "ports": "list[V1ContainerPort]",
    "resources": "V1ResourceRequirements", "working_dir": "str", "image_pull_policy": "str"})
V1PodSpec = _model("V1PodSpec", {
    "containers": "list[V1Container]", "restart_policy": "str", "scheduler_name": "str",
    "node_selector": "dict(str, str)", "volumes": "list[object]"})
V1PodTemplateSpec = _model("V1PodTemplateSpec", {"metadata": "V1ObjectMeta", "spec": "V1PodSpec"})
V1ListMeta = _model("V1ListMeta", {"resource_version": "str", "_continue": "str"},
                    {"_continue": "continue"})

# ---- common job models -----------------------------------------------------------------
V1Time = _model("V1Time", {})
V1JobCondition = _model("V1JobCondition", {
    "last_transition_time": "V1Time", "last_update_time": "V1Time", "message": "str",
    "reason": "str", "status": "str", "type": "str"})
V1ReplicaStatus = _model("V1ReplicaStatus", {"active": "int", "failed": "int", "succeeded": "int"})
V1JobStatus = _model("V1JobStatus", {
    "completion_time": "V1Time", "conditions": "list[V1JobCondition]",
    "last_reconcile_time": "V1Time", "replica_statuses": "dict(str, V1ReplicaStatus)",
    "start_time": "V1Time"})
V1ReplicaSpec = _model("V1ReplicaSpec", {"replicas": "int", "restart_policy": "str",
                                         "template": "V1PodTemplateSpec"})
V1SchedulingPolicy = _model("V1SchedulingPolicy", {
    "min_available": "int", "queue": "str", "min_resources": "dict(str, str)",
    "priority_class": "str"})
V1RunPolicy = _model("V1RunPolicy", {
    "clean_pod_policy": "str", "ttl_seconds_after_finished": "int",
    "active_deadline_seconds": "int", "backoff_limit": "int",
    "scheduling_policy": "V1SchedulingPolicy"})

_FLAT = {"active_deadline_seconds": "int", "backoff_limit": "int", "clean_pod_policy": "str",
         "ttl_seconds_after_finished": "int", "run_policy": "V1RunPolicy"}


def _job_models(prefix: str, kind: str, specs_attr: str, specs_key: str, extra=None):
    spec = _model(f"V1{prefix}JobSpec", {**_FLAT, specs_attr: "dict(str, V1ReplicaSpec)",
                                         **(extra or {})}, {specs_attr: specs_key})
    job = _model(f"V1{prefix}Job", {"api_version": "str", "kind": "str",
                                    "metadata": "V1ObjectMeta", "spec": spec.__name__,
                                    "status": "V1JobStatus"})
    lst = _model(f"V1{prefix}JobList", {"api_version": "str", "items": f"list[{job.__name__}]",
                                        "kind": "str", "metadata": "V1ListMeta"})
    return spec, job, lst


V1TFJobSpec, V1TFJob, V1TFJobList = _job_models(
    "TF", "TFJob", "tf_replica_specs", "tfReplicaSpecs",
    {"success_policy": "str", "enable_dynamic_worker": "bool"})
V1PyTorchJobSpec, V1PyTorchJob, V1PyTorchJobList = _job_models(
    "PyTorch", "PyTorchJob", "pytorch_replica_specs", "pytorchReplicaSpecs")
V1MXJobSpec, V1MXJob, V1MXJobList = _job_models(
    "MX", "MXJob", "mx_replica_specs", "mxReplicaSpecs", {"job_mode": "str"})
V1XGBoostJobSpec, V1XGBoostJob, V1XGBoostJobList = _job_models(
    "XGBoost", "XGBoostJob", "xgb_replica_specs", "xgbReplicaSpecs")
