"""TFJob API: constants, defaulting, validation.

Reference: pkg/apis/tensorflow/v1/{constants.go:21-37, types.go:48-94, defaults.go:38-115,
util.go:23-38}, pkg/apis/tensorflow/validation/validation.go:27-66.
"""
from __future__ import annotations

from typing import Any, Dict

from . import common as c

PS, WORKER, CHIEF, MASTER, EVALUATOR = "PS", "Worker", "Chief", "Master", "Evaluator"
SUCCESS_POLICY_DEFAULT = ""
SUCCESS_POLICY_ALL_WORKERS = "AllWorkers"

INFO = c.KindInfo(
    kind="TFJob", plural="tfjobs", singular="tfjob", specs_key="tfReplicaSpecs",
    container="tensorflow", port_name="tfjob-port", port=2222,
    restart_policy=c.RESTART_NEVER, clean_pod_policy=c.CLEAN_RUNNING,
    replica_types=(PS, WORKER, CHIEF, MASTER, EVALUATOR),
    normalised_types=(PS, WORKER, CHIEF, MASTER, EVALUATOR),
    controller_name="tf-operator")

# condition reasons (pkg/controller.v1/tensorflow/status.go:34-45)
REASON_CREATED = "TFJobCreated"
REASON_SUCCEEDED = "TFJobSucceeded"
REASON_RUNNING = "TFJobRunning"
REASON_FAILED = "TFJobFailed"
REASON_RESTARTING = "TFJobRestarting"


def is_chief_or_master(rtype: str) -> bool:
    return rtype in (CHIEF, MASTER)


def is_worker(rtype: str) -> bool:
    return rtype == WORKER


def is_evaluator(rtype: str) -> bool:
    return rtype == EVALUATOR


def set_defaults(job: Dict[str, Any]) -> Dict[str, Any]:
    """SetDefaults_TFJob: cleanPodPolicy Running, successPolicy "", canonical type names,
    replicas 1, restartPolicy Never, default port on EVERY replica type."""
    c.set_common_defaults(job, INFO)
    spec = job.setdefault("spec", {})
    if spec.get("successPolicy") is None:
        spec["successPolicy"] = SUCCESS_POLICY_DEFAULT
    return job


def validate(spec: Dict[str, Any]) -> None:
    """ValidateV1TFJobSpec."""
    specs = (spec or {}).get(INFO.specs_key)
    if specs is None:
        raise c.ValidationError("TFJobSpec is not valid")
    found_chief = 0
    for rtype, value in specs.items():
        containers = c.containers_of(value)
        if value is None or not containers:
            raise c.ValidationError(
                f"TFJobSpec is not valid: containers definition expected in {rtype}")
        if is_chief_or_master(rtype):
            found_chief += 1
        named = 0
        for ctr in containers:
            if not ctr.get("image"):
                raise c.ValidationError(
                    f"TFJobSpec is not valid: Image is undefined in the container of {rtype}")
            if ctr.get("name") == INFO.container:
                named += 1
        if named == 0:
            raise c.ValidationError(
                f"TFJobSpec is not valid: There is no container named {INFO.container} in {rtype}")
    if found_chief > 1:
        raise c.ValidationError("TFJobSpec is not valid: more than 1 chief/master found")
