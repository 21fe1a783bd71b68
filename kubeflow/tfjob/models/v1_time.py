"""sdk/python/kubeflow/tfjob/models/v1_time.py"""
from tf_operator_b200.sdk.models import V1Time  # noqa: F401
