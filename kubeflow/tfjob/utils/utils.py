"""sdk/python/kubeflow/tfjob/utils/utils.py:19-74"""
from tf_operator_b200.sdk.utils import *  # noqa: F401,F403
from tf_operator_b200.sdk.utils import (get_current_k8s_namespace, get_default_target_namespace,  # noqa: F401
                                        get_labels, is_running_in_k8s, set_job_namespace,
                                        to_selector)

set_tfjob_namespace = set_job_namespace
