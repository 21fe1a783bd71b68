"""MXJob API: constants, defaulting, validation.

Reference: pkg/apis/mxnet/v1/{constants.go:5-21, mxjob_types.go:23-79, defaults.go:71-110},
pkg/apis/mxnet/validation/validation.go:37-70.
"""
from __future__ import annotations

from typing import Any, Dict

from . import common as c

SCHEDULER, SERVER, WORKER = "Scheduler", "Server", "Worker"
TUNER_TRACKER, TUNER_SERVER, TUNER = "TunerTracker", "TunerServer", "Tuner"
MX_TRAIN, MX_TUNE = "MXTrain", "MXTune"
ANNOTATION_TUNER_SERVER_KEY = "tuner-server-key"

INFO = c.KindInfo(
    kind="MXJob", plural="mxjobs", singular="mxjob", specs_key="mxReplicaSpecs",
    container="mxnet", port_name="mxjob-port", port=9091,
    restart_policy=c.RESTART_NEVER, clean_pod_policy=c.CLEAN_ALL,
    replica_types=(SCHEDULER, SERVER, WORKER, TUNER_TRACKER, TUNER_SERVER, TUNER),
    # only these three are case-normalised (defaults.go:71-75)
    normalised_types=(SCHEDULER, SERVER, WORKER),
    controller_name="mxnet-operator")

REASON_CREATED = "MXJobCreated"
REASON_SUCCEEDED = "MXJobSucceeded"
REASON_RUNNING = "MXJobRunning"
REASON_FAILED = "MXJobFailed"
REASON_RESTARTING = "MXJobRestarting"


def set_defaults(job: Dict[str, Any]) -> Dict[str, Any]:
    """SetDefaults_MXJob: cleanPodPolicy All; default port on every replica type."""
    c.set_common_defaults(job, INFO)
    return job


def validate(spec: Dict[str, Any]) -> None:
    """ValidateV1MXJobSpec (every failure reads "MXJobSpec is not valid" except the
    scheduler-count one)."""
    specs = (spec or {}).get(INFO.specs_key)
    if specs is None:
        raise c.ValidationError("MXJobSpec is not valid")
    found_scheduler = 0
    for rtype, value in specs.items():
        containers = c.containers_of(value)
        if value is None or not containers:
            raise c.ValidationError("MXJobSpec is not valid")
        if rtype == SCHEDULER:
            found_scheduler += 1
        named = 0
        for ctr in containers:
            if not ctr.get("image"):
                raise c.ValidationError("MXJobSpec is not valid")
            if ctr.get("name") == INFO.container:
                named += 1
        if named == 0:
            raise c.ValidationError("MXJobSpec is not valid")
    if found_scheduler > 1:
        raise c.ValidationError("more than 1 scheduler found")
