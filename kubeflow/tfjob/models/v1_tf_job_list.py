"""sdk/python/kubeflow/tfjob/models/v1_tf_job_list.py"""
from tf_operator_b200.sdk.models import V1TFJobList  # noqa: F401
