"""sdk/python/kubeflow/tfjob/api/tf_job_watch.py:23-59"""
from tf_operator_b200.sdk.watch import watch  # noqa: F401
