"""XGBoostJob API: constants, defaulting, validation.

Reference: pkg/apis/xgboost/v1/{constants.go:20-35, xgboostjob_types.go:26-66,
defaults.go:93-110}, pkg/apis/xgboost/validation/validation.go:25-79.
"""
from __future__ import annotations

from typing import Any, Dict

from . import common as c
from .pytorch import _validate_master_worker

MASTER, WORKER = "Master", "Worker"

INFO = c.KindInfo(
    kind="XGBoostJob", plural="xgboostjobs", singular="xgboostjob",
    specs_key="xgbReplicaSpecs", container="xgboost", port_name="xgboostjob-port", port=9999,
    restart_policy=c.RESTART_NEVER, clean_pod_policy=c.CLEAN_ALL,
    replica_types=(MASTER, WORKER), normalised_types=(MASTER, WORKER),
    controller_name="xgboostjob-operator")

REASON_CREATED = "XGBoostJobCreated"
REASON_SUCCEEDED = "XGBoostJobSucceeded"
REASON_RUNNING = "XGBoostJobRunning"
REASON_FAILED = "XGBoostJobFailed"
REASON_RESTARTING = "XGBoostJobRestarting"


def set_defaults(job: Dict[str, Any]) -> Dict[str, Any]:
    """SetDefaults_XGBoostJob: cleanPodPolicy All; default port on every replica type."""
    c.set_common_defaults(job, INFO)
    return job


def validate(spec: Dict[str, Any]) -> None:
    """ValidateXGBoostJobSpec — same rules as PyTorchJob with xgboost names."""
    _validate_master_worker(spec, INFO, "XGBoostJobSpec", "XGBoostReplicaType",
                            "XGBoostReplicaType")
