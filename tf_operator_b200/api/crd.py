"""CRD manifests for the four kinds (`python -m tf_operator_b200.api.crd manifests/base`).

Same identity and structure as the reference's controller-gen output
(manifests/base/kubeflow.org_*.yaml: apiextensions.k8s.io/v1, group kubeflow.org, version v1
served+storage, Namespaced, ``subresources.status``, required spec fields, runPolicy and
status schemas from :36-108 and :6816-6896 of kubeflow.org_tfjobs.yaml), including the full
structural ``PodTemplateSpec`` (:109-6815) so a structural apiserver prunes and validates the pod
template exactly as it does under the reference's CRDs.

The STRUCTURE (fields, types, formats, required lists) is generated here from the API model. Two
data files under ``api/schema/`` complete it: ``core_v1_podtemplatespec.json.gz`` - the
core/v1 PodTemplateSpec schema as controller-gen v0.4.1 emits it, identical for the four kinds
and therefore stored once - and ``crd_descriptions.json``, the field documentation strings keyed
by kind and schema path. ``yaml.safe_load`` of every emitted manifest equals ``yaml.safe_load``
of the reference's (tests/test_manifests.py).
"""
from __future__ import annotations

import copy
import gzip
import json
import sys
from functools import lru_cache
from pathlib import Path

import yaml

from . import KINDS

_INT32 = {"type": "integer", "format": "int32"}
_INT64 = {"type": "integer", "format": "int64"}
_TIME = {"type": "string", "format": "date-time"}

RUN_POLICY = {
    "description": "RunPolicy encapsulates various runtime policies of the distributed training job, "
                   "for example how to clean up resources and how long the job can stay active.",
    "type": "object",
    "properties": {
        "activeDeadlineSeconds": {**_INT64, "description": "Specifies the duration in seconds relative to "
                                  "the startTime that the job may be active before the system tries to terminate it; "
                                  "value must be positive integer."},
        "backoffLimit": {**_INT32, "description": "Optional number of retries before marking this job failed."},
        "cleanPodPolicy": {"type": "string", "description": "CleanPodPolicy defines the policy to kill pods "
                           "after the job completes. Default to Running."},
        "schedulingPolicy": {
            "description": "SchedulingPolicy defines the policy related to scheduling, e.g. gang-scheduling",
            "type": "object",
            "properties": {
                "minAvailable": _INT32,
                "minResources": {"type": "object", "additionalProperties": {
                    "anyOf": [{"type": "integer"}, {"type": "string"}],
                    "pattern": "^(\\+|-)?(([0-9]+(\\.[0-9]*)?)|(\\.[0-9]+))(([KMGTPE]i)|[numkMGTPE]|([eE](\\+|-)?(([0-9]+(\\.[0-9]*)?)|(\\.[0-9]+))))?$",
                    "x-kubernetes-int-or-string": True}},
                "priorityClass": {"type": "string"},
                "queue": {"type": "string"},
            },
        },
        "ttlSecondsAfterFinished": {**_INT32, "description": "TTLSecondsAfterFinished is the TTL to clean up "
                                    "jobs. It may take extra ReconcilePeriod seconds for the cleanup, since reconcile "
                                    "gets called periodically. Default to infinite."},
    },
}

REPLICA_SPEC = {
    "description": "ReplicaSpec is a description of the replica",
    "type": "object",
    "properties": {
        "replicas": {**_INT32, "description": "Replicas is the desired number of replicas of the given "
                     "template. If unspecified, defaults to 1."},
        "restartPolicy": {"type": "string", "description": "Restart policy for all replicas within the job. "
                          "One of Always, OnFailure, Never and ExitCode. Default to Never."},
        "template": {"description": "Template is the object that describes the pod that will be created for "
                     "this replica. RestartPolicy in PodTemplateSpec will be overide by RestartPolicy in ReplicaSpec",
                     "type": "object", "x-kubernetes-preserve-unknown-fields": True},
    },
}

STATUS = {
    "description": "JobStatus represents the current observed state of the training Job.",
    "type": "object",
    "required": ["conditions", "replicaStatuses"],
    "properties": {
        "completionTime": _TIME,
        "conditions": {"type": "array", "items": {
            "description": "JobCondition describes the state of the job at a certain point.",
            "type": "object", "required": ["status", "type"],
            "properties": {"lastTransitionTime": _TIME, "lastUpdateTime": _TIME,
                           "message": {"type": "string"}, "reason": {"type": "string"},
                           "status": {"type": "string"}, "type": {"type": "string"}}}},
        "lastReconcileTime": _TIME,
        "replicaStatuses": {"type": "object", "additionalProperties": {
            "description": "ReplicaStatus represents the current observed state of the replica.",
            "type": "object",
            "properties": {"active": _INT32, "failed": _INT32, "succeeded": _INT32}}},
        "startTime": _TIME,
    },
}


_SCHEMA_DIR = Path(__file__).resolve().parent / "schema"


@lru_cache(maxsize=1)
def pod_template_schema() -> dict:
    """core/v1 PodTemplateSpec as an OpenAPI v3 structural schema (data file, one copy)."""
    with gzip.open(_SCHEMA_DIR / "core_v1_podtemplatespec.json.gz", "rt") as f:
        return json.load(f)


@lru_cache(maxsize=1)
def _descriptions() -> dict:
    return json.loads((_SCHEMA_DIR / "crd_descriptions.json").read_text())


def _strip_descriptions(node):
    if isinstance(node, dict):
        return {k: _strip_descriptions(v) for k, v in node.items()
                if not (k == "description" and isinstance(v, str))}
    if isinstance(node, list):
        return [_strip_descriptions(v) for v in node]
    return node


def _document(schema: dict, kind: str) -> dict:
    """Attach the per-kind field documentation (schema path -> text) to a bare structure."""
    for path, text in _descriptions()[kind].items():
        node = schema
        for key in (path.split("/") if path else []):
            node = node[key]
        node["description"] = text
    return schema


def crd_for(mod, full: bool = True) -> dict:
    """The CustomResourceDefinition of one kind. ``full=False`` keeps the pod template open
    (``x-kubernetes-preserve-unknown-fields``) and the generator's own short descriptions - the
    compact form the in-process API server registers."""
    crd = _crd_structure(mod)
    if not full:
        return crd
    i = mod.INFO
    schema = _strip_descriptions(crd["spec"]["versions"][0]["schema"]["openAPIV3Schema"])
    rs = schema["properties"]["spec"]["properties"][i.specs_key]["additionalProperties"]
    rs["properties"]["template"] = {"type": "object"}
    schema = _document(schema, i.kind)
    rs = schema["properties"]["spec"]["properties"][i.specs_key]["additionalProperties"]
    rs["properties"]["template"] = copy.deepcopy(pod_template_schema())
    crd["spec"]["versions"][0]["schema"]["openAPIV3Schema"] = schema
    # controller-gen leaves these behind; kept so the documents compare equal
    crd["metadata"]["creationTimestamp"] = None
    crd["status"] = {"acceptedNames": {"kind": "", "plural": ""}, "conditions": [],
                     "storedVersions": []}
    return crd


def _crd_structure(mod) -> dict:
    i = mod.INFO
    spec_props = {"runPolicy": RUN_POLICY,
                  i.specs_key: {"type": "object", "additionalProperties": REPLICA_SPEC}}
    required = [i.specs_key]
    if i.kind == "TFJob":
        spec_props["successPolicy"] = {"type": "string"}
        spec_props["enableDynamicWorker"] = {"type": "boolean"}
    if i.kind == "MXJob":
        spec_props["jobMode"] = {"type": "string"}
        required = ["jobMode", i.specs_key]
    return {
        "apiVersion": "apiextensions.k8s.io/v1", "kind": "CustomResourceDefinition",
        "metadata": {"name": f"{i.plural}.kubeflow.org",
                     "annotations": {"controller-gen.kubebuilder.io/version": "v0.4.1"}},
        "spec": {
            "group": "kubeflow.org",
            "names": {"kind": i.kind, "listKind": i.kind + "List", "plural": i.plural,
                      "singular": i.singular},
            "scope": "Namespaced",
            "versions": [{
                "name": "v1", "served": True, "storage": True, "subresources": {"status": {}},
                "schema": {"openAPIV3Schema": {
                    "description": f"{i.kind} is the Schema for the {i.plural} API",
                    "type": "object",
                    "properties": {
                        "apiVersion": {"type": "string"}, "kind": {"type": "string"},
                        "metadata": {"type": "object"},
                        "spec": {"type": "object", "required": sorted(required),
                                 "properties": dict(sorted(spec_props.items()))},
                        "status": STATUS,
                    }}},
            }],
        },
    }


def write_all(out_dir: str) -> list:
    out = Path(out_dir)
    out.mkdir(parents=True, exist_ok=True)
    paths = []
    for mod in KINDS.values():
        p = out / f"kubeflow.org_{mod.INFO.plural}.yaml"
        p.write_text("\n---\n" + yaml.safe_dump(crd_for(mod), sort_keys=True, width=80))
        paths.append(p)
    return paths


if __name__ == "__main__":
    for p in write_all(sys.argv[1] if len(sys.argv) > 1 else "manifests/base"):
        print(p)
