"""Common reconcile entry shared by the four per-kind reconcilers.

Shape of ``Reconcile`` in the reference (pkg/controller.v1/tensorflow/tfjob_controller.go:118-159):
Get -> validate (log only, does NOT abort) -> deletion / expectation gate -> Scheme.Default ->
ReconcileJobs. ``on_owner_create`` is the create predicate (defaults + "Created" condition,
:837-854).
"""
from __future__ import annotations

import copy
import logging
from typing import Any, Dict, Optional

from ..api import common as c
from ..engine import expectations as exp
from ..engine.job_controller import JobController
from ..store.memory import ApiError

log = logging.getLogger("tf_operator_b200.kinds")


class KindReconciler(JobController):
    created_reason = "JobCreated"
    created_message = "{kind} {name} is created."
    # legacy TFController marks un-parsable specs Failed (job.go:62-112); the new reconcilers
    # only log. Off by default = new-manager behaviour.
    fail_invalid_spec = False
    invalid_spec_reason = "InvalidJobSpec"

    # ---- create predicate ----------------------------------------------------
    def on_owner_create(self, job: Dict[str, Any]) -> Dict[str, Any]:
        """Defaults + Created condition; persists the status if a store is attached."""
        self.module.set_defaults(job)
        status = c.status_of(job)
        msg = self.created_message.format(kind=self.info.kind, name=c.name_of(job))
        before = c.deepcopy(status)
        c.update_job_conditions(status, c.JOB_CREATED, self.created_reason, msg)
        if self.metrics is not None and before != status:
            self.metrics.jobs_created.labels(c.namespace_of(job)).inc()
        if self.store is not None and before != status:
            try:
                self.update_job_status_in_api_server(job, status)
            except ApiError:
                pass
        return job

    # ---- Reconcile -------------------------------------------------------------
    def reconcile(self, namespace: str, name: str) -> Optional[Dict[str, Any]]:
        try:
            job = self.store.get(self.info.plural, namespace, name)
        except ApiError as e:
            if e.code == 404:
                log.info("unable to fetch %s %s/%s (deleted)", self.info.kind, namespace, name)
                return None
            raise
        return self.reconcile_object(job)

    def reconcile_object(self, job: Dict[str, Any]) -> Optional[Dict[str, Any]]:
        try:
            self.module.validate(job.get("spec") or {})
        except c.ValidationError as e:
            log.info("%s failed validation: %s", c.key_of(job), e)
            if self.fail_invalid_spec:
                self._mark_invalid(job, str(e))
                return job
        job_key = c.key_of(job)
        specs = c.replica_specs(job, self.info)
        need = exp.satisfied_expectations(self.expectations, job_key, list(specs.keys()))
        # elastic jobs are always synced, stale expectations or not (controller.go:336)
        if (job.get("spec") or {}).get("enableDynamicWorker"):
            need = True
        if not need or job.get("metadata", {}).get("deletionTimestamp"):
            log.info("reconcile cancelled: job %s does not need a reconcile or is being deleted "
                     "(sync=%s)", job_key, need)
            return job
        self.module.set_defaults(job)
        status = c.status_of(job)
        if not status["conditions"]:
            self.on_owner_create(job)
            status = c.status_of(job)
        specs = c.replica_specs(job, self.info)
        self.reconcile_jobs(job, specs, status, self.run_policy_for(job))
        return job

    def _mark_invalid(self, job, message: str) -> None:
        status = c.status_of(job)
        msg = f"Failed to marshal the object to {self.info.kind}; the spec is invalid: {message}"
        self.recorder.event(job, "Warning", self.invalid_spec_reason, msg)
        if status.get("completionTime") is None:
            status["completionTime"] = c.now_rfc3339()
        c.update_job_conditions(status, c.JOB_FAILED, self.invalid_spec_reason, msg)
        self.update_job_status_in_api_server(job, status)

    # ---- dependents' watch predicates (pkg/common/util/reconciler.go:38-157) -----
    def on_dependent_create(self, resource: str, obj) -> Optional[str]:
        """Returns the owning job key to enqueue, or None to drop the event."""
        rtype = (obj.get("metadata", {}).get("labels") or {}).get(c.LABEL_REPLICA_TYPE)
        if not rtype:
            return None
        ref = c.get_controller_of(obj)
        if ref is None or ref.get("kind") != self.info.kind:
            return None
        job_key = f"{c.namespace_of(obj)}/{ref['name']}"
        key = (exp.gen_expectation_pods_key if resource == "pods"
               else exp.gen_expectation_services_key)(job_key, rtype)
        self.expectations.creation_observed(key)
        return job_key

    def on_dependent_delete(self, resource: str, obj) -> Optional[str]:
        rtype = (obj.get("metadata", {}).get("labels") or {}).get(c.LABEL_REPLICA_TYPE)
        if not rtype:
            return None
        ref = c.get_controller_of(obj)
        if ref is None or ref.get("kind") != self.info.kind:
            return None
        job_key = f"{c.namespace_of(obj)}/{ref['name']}"
        key = (exp.gen_expectation_pods_key if resource == "pods"
               else exp.gen_expectation_services_key)(job_key, rtype)
        self.expectations.deletion_observed(key)
        return job_key

    def on_dependent_update(self, resource: str, old, new) -> Optional[str]:
        """Ignore same-resourceVersion resyncs; resolve the owner by Kind+Name+UID."""
        if old is not None and old.get("metadata", {}).get("resourceVersion") == \
                new.get("metadata", {}).get("resourceVersion"):
            return None
        ref = c.get_controller_of(new)
        if ref is None or ref.get("kind") != self.info.kind:
            return None
        if self.store is not None:
            try:
                job = self.store.get(self.info.plural, c.namespace_of(new), ref["name"])
            except ApiError:
                return None
            if job.get("metadata", {}).get("uid") != ref.get("uid"):
                return None
        return f"{c.namespace_of(new)}/{ref['name']}"
