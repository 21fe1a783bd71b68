"""SDK helpers (sdk/python/kubeflow/tfjob/utils/utils.py:19-74)."""
import os

from . import constants


def is_running_in_k8s() -> bool:
    return os.path.isdir("/var/run/secrets/kubernetes.io/")


def get_current_k8s_namespace() -> str:
    with open("/var/run/secrets/kubernetes.io/serviceaccount/namespace", "r") as f:
        return f.readline()


def get_default_target_namespace() -> str:
    if not is_running_in_k8s():
        return os.environ.get("TFOP_POD_NAMESPACE", "default")
    return get_current_k8s_namespace()


def set_job_namespace(job) -> str:
    """Namespace of a job given as a model object (attribute access) or a dict."""
    if isinstance(job, dict):
        ns = (job.get("metadata") or {}).get("namespace")
    else:
        md = getattr(job, "metadata", None)
        ns = getattr(md, "namespace", None) if md is not None and not isinstance(md, dict) \
            else (md or {}).get("namespace")
    return ns or get_default_target_namespace()


set_tfjob_namespace = set_job_namespace


def get_labels(name, master=False, replica_type=None, replica_index=None):
    """Label set used to select a job's pods."""
    labels = {constants.TFJOB_GROUP_LABEL: "kubeflow.org", constants.TFJOB_NAME_LABEL: name}
    if master:
        labels[constants.TFJOB_ROLE_LABEL] = "master"
    if replica_type:
        labels[constants.TFJOB_TYPE_LABEL] = str.lower(replica_type)
    if replica_index is not None:
        labels[constants.TFJOB_INDEX_LABEL] = str(replica_index)
    return labels


def to_selector(labels) -> str:
    return ",".join(f"{k}={v}" for k, v in labels.items())
