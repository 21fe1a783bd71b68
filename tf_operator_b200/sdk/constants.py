"""SDK constants (sdk/python/kubeflow/tfjob/constants/constants.py:17-33)."""
import os

TFJOB_GROUP = "kubeflow.org"
TFJOB_KIND = "TFJob"
TFJOB_PLURAL = "tfjobs"
TFJOB_VERSION = os.environ.get("TFJOB_VERSION", "v1")
TFJOB_LOGLEVEL = os.environ.get("TFJOB_LOGLEVEL", "INFO").upper()

PYTORCHJOB_GROUP = "kubeflow.org"
PYTORCHJOB_KIND = "PyTorchJob"
PYTORCHJOB_PLURAL = "pytorchjobs"
PYTORCHJOB_VERSION = os.environ.get("PYTORCHJOB_VERSION", "v1")

# How long to wait in seconds for requests to the ApiServer
APISERVER_TIMEOUT = 120

# TFJob Labels Name
TFJOB_GROUP_LABEL = "group-name"
TFJOB_NAME_LABEL = "job-name"
TFJOB_TYPE_LABEL = "replica-type"
TFJOB_INDEX_LABEL = "replica-index"
TFJOB_ROLE_LABEL = "job-role"
