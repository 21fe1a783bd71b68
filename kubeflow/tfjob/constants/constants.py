"""sdk/python/kubeflow/tfjob/constants/constants.py:17-33"""
from tf_operator_b200.sdk.constants import *  # noqa: F401,F403
