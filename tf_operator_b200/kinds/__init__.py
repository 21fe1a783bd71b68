"""Per-framework reconcilers (pkg/controller.v1/{tensorflow,pytorch,mxnet,xgboost})."""
from .mxnet import MXJobReconciler
from .pytorch import PyTorchJobReconciler
from .tf import TFJobReconciler
from .xgboost import XGBoostJobReconciler

RECONCILERS = {
    "TFJob": TFJobReconciler, "PyTorchJob": PyTorchJobReconciler,
    "MXJob": MXJobReconciler, "XGBoostJob": XGBoostJobReconciler,
}
