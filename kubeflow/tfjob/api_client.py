"""sdk/python/kubeflow/tfjob/api_client.py"""
from tf_operator_b200.sdk.api_client import ApiClient  # noqa: F401
