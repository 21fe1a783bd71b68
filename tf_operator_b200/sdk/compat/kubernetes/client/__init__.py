"""``kubernetes.client`` models used by reference-SDK scripts (sdk/python/test/test_e2e.py:15-19,
sdk/python/examples/kubeflow-tfjob-sdk.ipynb), served by tf_operator_b200.sdk.models."""
from tf_operator_b200.sdk.models import (V1Container, V1ContainerPort, V1EnvVar, V1ObjectMeta,  # noqa: F401
                                         V1PodSpec, V1PodTemplateSpec, V1ResourceRequirements)


class V1DeleteOptions(dict):
    """Accepted and ignored by JobClient.delete (tf_job_client.py:199-220 passes one)."""
