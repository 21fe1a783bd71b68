"""sdk/python/kubeflow/tfjob/models/v1_replica_status.py"""
from tf_operator_b200.sdk.models import V1ReplicaStatus  # noqa: F401
