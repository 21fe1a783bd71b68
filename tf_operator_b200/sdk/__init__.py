"""Python SDK: kubeflow.tfjob-compatible TFJobClient plus PyTorchJobClient / MXJobClient /
XGBoostJobClient (sdk/python/kubeflow/tfjob/__init__.py:17-36 for the exported names)."""
from . import constants, utils
from .api_client import ApiClient
from .configuration import Configuration
from .rest import ApiException
from .client import JobClient, MXJobClient, PyTorchJobClient, TFJobClient, XGBoostJobClient
from .models import (V1Container, V1ContainerPort, V1EnvVar, V1JobCondition, V1JobStatus,
                     V1MXJob, V1MXJobList, V1MXJobSpec, V1ObjectMeta, V1PodSpec,
                     V1PodTemplateSpec, V1PyTorchJob, V1PyTorchJobList, V1PyTorchJobSpec,
                     V1ReplicaSpec, V1ReplicaStatus, V1ResourceRequirements, V1RunPolicy,
                     V1SchedulingPolicy, V1TFJob, V1TFJobList, V1TFJobSpec, V1Time,
                     V1XGBoostJob, V1XGBoostJobList, V1XGBoostJobSpec)
from .watch import watch
