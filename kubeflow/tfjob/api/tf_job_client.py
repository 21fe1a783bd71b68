"""sdk/python/kubeflow/tfjob/api/tf_job_client.py:54-442"""
from tf_operator_b200.sdk.client import TFJobClient  # noqa: F401
