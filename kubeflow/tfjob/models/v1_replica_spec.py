"""sdk/python/kubeflow/tfjob/models/v1_replica_spec.py"""
from tf_operator_b200.sdk.models import V1ReplicaSpec  # noqa: F401
