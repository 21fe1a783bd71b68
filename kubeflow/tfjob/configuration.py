"""sdk/python/kubeflow/tfjob/configuration.py"""
from tf_operator_b200.sdk.configuration import Configuration  # noqa: F401
