"""sdk/python/kubeflow/tfjob/models/v1_tf_job_spec.py"""
from tf_operator_b200.sdk.models import V1TFJobSpec  # noqa: F401
