"""Prometheus metrics with the reference's names:
tf_operator_jobs_{created,deleted,successful,failed,restarted}_total{job_namespace} and the
tf_operator_is_leader gauge (pkg/controller.v1/tensorflow/job.go:29-37, controller.go:70-76,
status.go:47-62, pod.go:57-65; cmd/tf-operator.v1/app/server.go:64-69)."""
from __future__ import annotations

from prometheus_client import CollectorRegistry, Counter, Gauge, Histogram, generate_latest


class OperatorMetrics:
    def __init__(self, registry: CollectorRegistry | None = None):
        self.registry = registry or CollectorRegistry()
        mk = lambda n, d: Counter(n, d, ["job_namespace"], registry=self.registry)  # noqa: E731
        # prometheus_client appends _total itself
        self.jobs_created = mk("tf_operator_jobs_created", "Counts number of TF jobs created")
        self.jobs_deleted = mk("tf_operator_jobs_deleted", "Counts number of TF jobs deleted")
        self.jobs_successful = mk("tf_operator_jobs_successful", "Counts number of TF jobs successful")
        self.jobs_failed = mk("tf_operator_jobs_failed", "Counts number of TF jobs failed")
        self.jobs_restarted = mk("tf_operator_jobs_restarted", "Counts number of TF jobs restarted")
        self.is_leader = Gauge("tf_operator_is_leader", "Is this client the leader of this "
                               "tf-operator client set?", registry=self.registry)
        self.reconcile_seconds = Histogram(
            "training_operator_reconcile_seconds", "Wall time of one reconcile", ["kind"],
            registry=self.registry,
            buckets=(.0005, .001, .0025, .005, .01, .025, .05, .1, .25, .5, 1, 2.5, 5))
        self.launch_to_running_seconds = Histogram(
            "training_operator_launch_to_running_seconds",
            "Job creation to Running condition", ["kind"], registry=self.registry,
            buckets=(.005, .01, .025, .05, .1, .25, .5, 1, 2.5, 5, 10, 30, 60))

    def render(self) -> bytes:
        return generate_latest(self.registry)
