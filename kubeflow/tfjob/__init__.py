"""``kubeflow.tfjob``: the import path of the reference SDK (``pip install kubeflow-tfjob`` ->
``from kubeflow.tfjob import TFJobClient, V1TFJob, ...``,
sdk/python/kubeflow/tfjob/__init__.py:17-36), served by tf_operator_b200.sdk. Code written
against the reference SDK imports and runs unchanged; the client talks to the single-box
API server (in process, or the REST facade named by ``$TFOP_APISERVER``) instead of a
kube-apiserver."""
from __future__ import absolute_import

# utils and constants
from kubeflow.tfjob.utils import utils
from kubeflow.tfjob.constants import constants

# ApiClient
from kubeflow.tfjob.api_client import ApiClient
from kubeflow.tfjob.configuration import Configuration
from kubeflow.tfjob.api.tf_job_client import TFJobClient

# models
from kubeflow.tfjob.models.v1_job_condition import V1JobCondition
from kubeflow.tfjob.models.v1_job_status import V1JobStatus
from kubeflow.tfjob.models.v1_replica_spec import V1ReplicaSpec
from kubeflow.tfjob.models.v1_replica_status import V1ReplicaStatus
from kubeflow.tfjob.models.v1_tf_job import V1TFJob
from kubeflow.tfjob.models.v1_tf_job_list import V1TFJobList
from kubeflow.tfjob.models.v1_tf_job_spec import V1TFJobSpec
from kubeflow.tfjob.models.v1_time import V1Time

__version__ = "0.1.4"
