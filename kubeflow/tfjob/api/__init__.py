from kubeflow.tfjob.api.tf_job_client import TFJobClient  # noqa: F401
