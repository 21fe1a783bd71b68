"""``kubeflow`` namespace package (other kubeflow-* distributions may add to it)."""
__path__ = __import__("pkgutil").extend_path(__path__, __name__)
