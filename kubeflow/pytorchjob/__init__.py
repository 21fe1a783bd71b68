"""``kubeflow.pytorchjob``: PyTorchJobClient and models under the naming scheme of
``kubeflow.tfjob``. The reference ships only the TFJob SDK
(sdk/python/kubeflow/tfjob/__init__.py:25); this is its PyTorchJob analogue."""
from tf_operator_b200.sdk import constants, utils  # noqa: F401
from tf_operator_b200.sdk.api_client import ApiClient  # noqa: F401
from tf_operator_b200.sdk.configuration import Configuration  # noqa: F401
from tf_operator_b200.sdk.client import PyTorchJobClient  # noqa: F401
from tf_operator_b200.sdk.models import (V1JobCondition, V1JobStatus, V1PyTorchJob,  # noqa: F401
                                         V1PyTorchJobList, V1PyTorchJobSpec, V1ReplicaSpec,
                                         V1ReplicaStatus, V1RunPolicy, V1Time)
