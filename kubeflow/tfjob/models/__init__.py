"""sdk/python/kubeflow/tfjob/models/__init__.py"""
from kubeflow.tfjob.models.v1_job_condition import V1JobCondition  # noqa: F401
from kubeflow.tfjob.models.v1_job_status import V1JobStatus  # noqa: F401
from kubeflow.tfjob.models.v1_replica_spec import V1ReplicaSpec  # noqa: F401
from kubeflow.tfjob.models.v1_replica_status import V1ReplicaStatus  # noqa: F401
from kubeflow.tfjob.models.v1_tf_job import V1TFJob  # noqa: F401
from kubeflow.tfjob.models.v1_tf_job_list import V1TFJobList  # noqa: F401
from kubeflow.tfjob.models.v1_tf_job_spec import V1TFJobSpec  # noqa: F401
from kubeflow.tfjob.models.v1_time import V1Time  # noqa: F401
