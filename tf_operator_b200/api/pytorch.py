"""PyTorchJob API: constants, defaulting, validation.

Reference: pkg/apis/pytorch/v1/{constants.go:21-37, pytorchjob_types.go:44-66,
defaults.go:88-106}, pkg/apis/pytorch/validation/validation.go:25-79.
"""
from __future__ import annotations

from typing import Any, Dict

from . import common as c

MASTER, WORKER = "Master", "Worker"

INFO = c.KindInfo(
    kind="PyTorchJob", plural="pytorchjobs", singular="pytorchjob",
    specs_key="pytorchReplicaSpecs", container="pytorch", port_name="pytorchjob-port",
    port=23456, restart_policy=c.RESTART_ON_FAILURE, clean_pod_policy=c.CLEAN_NONE,
    replica_types=(MASTER, WORKER), normalised_types=(MASTER, WORKER),
    controller_name="pytorchjob-operator")

REASON_CREATED = "PyTorchJobCreated"


def set_defaults(job: Dict[str, Any]) -> Dict[str, Any]:
    """SetDefaults_PyTorchJob: cleanPodPolicy None; default port only on Master."""
    c.set_common_defaults(job, INFO, port_types=(MASTER,))
    return job


def _validate_master_worker(spec, info: c.KindInfo, spec_name: str, type_name: str,
                            not_valid_prefix: str) -> None:
    specs = (spec or {}).get(info.specs_key)
    if specs is None:
        raise c.ValidationError(f"{spec_name} is not valid")
    master_exists = False
    valid = list(info.replica_types)
    for rtype, value in specs.items():
        containers = c.containers_of(value)
        if value is None or not containers:
            raise c.ValidationError(
                f"{spec_name} is not valid: containers definition expected in {rtype}")
        if rtype not in valid:
            raise c.ValidationError(
                f"{type_name} is {rtype} but must be one of [{' '.join(valid)}]")
        default_present = False
        for ctr in containers:
            if not ctr.get("image"):
                raise c.ValidationError(
                    f"{not_valid_prefix} is not valid: Image is undefined in the container of {rtype}")
            if ctr.get("name") == info.container:
                default_present = True
        if not default_present:
            raise c.ValidationError(
                f"{not_valid_prefix} is not valid: There is no container named {info.container} in {rtype}")
        if rtype == "Master":
            master_exists = True
            reps = value.get("replicas")
            if reps is not None and int(reps) != 1:
                raise c.ValidationError(
                    f"{not_valid_prefix} is not valid: There must be only 1 master replica")
    if not master_exists:
        raise c.ValidationError(
            f"{not_valid_prefix} is not valid: Master ReplicaSpec must be present")


def validate(spec: Dict[str, Any]) -> None:
    """ValidateV1PyTorchJobSpec."""
    _validate_master_worker(spec, INFO, "PyTorchJobSpec", "PyTorchReplicaType", "PyTorchJobSpec")
