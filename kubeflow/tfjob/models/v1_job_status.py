"""sdk/python/kubeflow/tfjob/models/v1_job_status.py"""
from tf_operator_b200.sdk.models import V1JobStatus  # noqa: F401
