"""sdk/python/kubeflow/tfjob/models/v1_tf_job.py"""
from tf_operator_b200.sdk.models import V1TFJob  # noqa: F401
