"""kubeflow.org/v1 job APIs: TFJob, PyTorchJob, MXJob, XGBoostJob (PROJECT:5-38)."""
from . import common, mxnet, pytorch, tensorflow, xgboost

KINDS = {
    "TFJob": tensorflow, "PyTorchJob": pytorch, "MXJob": mxnet, "XGBoostJob": xgboost,
}
BY_PLURAL = {m.INFO.plural: m for m in KINDS.values()}


def module_for(kind_or_plural: str):
    if kind_or_plural in KINDS:
        return KINDS[kind_or_plural]
    low = kind_or_plural.lower()
    for k, m in KINDS.items():
        if low in (k.lower(), m.INFO.plural, m.INFO.singular):
            return m
    raise KeyError(f"unknown job kind {kind_or_plural!r}")


def set_defaults(job):
    return module_for(job["kind"]).set_defaults(job)


def validate(job):
    return module_for(job["kind"]).validate(job.get("spec") or {})
