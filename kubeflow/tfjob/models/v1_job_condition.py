"""sdk/python/kubeflow/tfjob/models/v1_job_condition.py"""
from tf_operator_b200.sdk.models import V1JobCondition  # noqa: F401
