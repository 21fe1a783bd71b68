"""sdk/python/kubeflow/tfjob/rest.py"""
from tf_operator_b200.sdk.rest import *  # noqa: F401,F403
from tf_operator_b200.sdk.rest import ApiException  # noqa: F401
