"""The shared reconcile engine: what ``kubeflow/common`` v0.3.4's JobController does for
the reference (the library is absent from the reference tree; SURVEY.md 2.4 lists the
contract and the in-tree tests that pin each rule).

    ReconcileJobs -> (terminal? clean up) | (limits exceeded? fail + clean up) |
                     (gang? SyncPodGroup) ; per replica type ReconcilePods + ReconcileServices ;
                     UpdateJobStatus ; write status iff changed

A per-kind controller (kinds/*.py) subclasses JobController and supplies
``set_cluster_spec``, ``is_master_role``, ``update_job_status`` (and, for TFJob, its own
``reconcile_pods``), mirroring the 17-method ControllerInterface of the reference
(pkg/controller.v1/tensorflow/tfjob_controller.go:205-640).
"""
from __future__ import annotations

import copy
import logging
import time
from dataclasses import dataclass
from typing import Any, Dict, List, Optional

from ..api import common as c
from ..store.memory import ApiError, ObjectStore
from . import expectations as exp
from .control import (FakePodControl, FakeServiceControl, RealPodControl, RealServiceControl,
                      Recorder)
from .workqueue import FakeWorkQueue, RateLimitingQueue

log = logging.getLogger("tf_operator_b200.engine")

REASON_EXITED_WITH_CODE = "ExitedWithCode"
REASON_POD_TEMPLATE_RESTART_POLICY = "SettedPodTemplateRestartPolicy"
REASON_POD_TEMPLATE_SCHEDULER_NAME = "SettedPodTemplateSchedulerName"
REASON_FAILED_DELETE_JOB = "FailedDeleteJob"
REASON_SUCCESSFUL_DELETE_JOB = "SuccessfulDeleteJob"


@dataclass
class JobControllerConfig:
    reconciler_sync_loop_period: float = 15.0   # controller.go:128
    enable_gang_scheduling: bool = False
    gang_scheduler_name: str = c.GANG_SCHEDULER_NAME


class JobController:
    """Engine + the per-kind hooks (override in kinds/*)."""

    module = None            # api module of the kind (api.tensorflow, ...)

    def __init__(self, store: Optional[ObjectStore], *, config: Optional[JobControllerConfig] = None,
                 pod_control=None, service_control=None, recorder: Optional[Recorder] = None,
                 workqueue=None, metrics=None):
        self.store = store
        self.info: c.KindInfo = self.module.INFO
        self.config = config or JobControllerConfig()
        self.recorder = recorder or Recorder(store, self.info.controller_name)
        self.pod_control = pod_control or RealPodControl(store, self.recorder)
        self.service_control = service_control or RealServiceControl(store, self.recorder)
        self.expectations = exp.ControllerExpectations()
        self.workqueue = workqueue if workqueue is not None else FakeWorkQueue()
        self.metrics = metrics
        self._clock = time.time

    # ------------------------------------------------------------------ identity
    def controller_name(self) -> str:
        return self.info.controller_name

    def get_default_container_name(self) -> str:
        return self.info.container

    def get_default_container_port_name(self) -> str:
        return self.info.port_name

    def gen_labels(self, job_name: str) -> Dict[str, str]:
        return c.gen_labels(job_name)

    def gen_owner_reference(self, job) -> Dict[str, Any]:
        return c.gen_owner_reference(job)

    # ------------------------------------------------------------------ hooks
    def set_cluster_spec(self, job, pod_template, rtype: str, index: str) -> None:
        raise NotImplementedError

    def is_master_role(self, replicas, rtype: str, index: int) -> bool:
        raise NotImplementedError

    def update_job_status(self, job, replicas, status) -> None:
        raise NotImplementedError

    def services_for(self, rtype: str) -> bool:
        """Whether replicas of this type get a headless Service."""
        return True

    def run_policy_for(self, job) -> Dict[str, Any]:
        return c.run_policy(job)

    # ------------------------------------------------------------------ reads
    def get_pods_for_job(self, job) -> List[Dict[str, Any]]:
        """List by GenLabels in the job's namespace, then claim: adopt matching orphans,
        release owned pods whose labels no longer match
        (tfjob_controller.go:251-289 via ControllerRefManager.ClaimPods)."""
        return self._claim("pods", job)

    def get_services_for_job(self, job) -> List[Dict[str, Any]]:
        return self._claim("services", job)

    def _list_candidates(self, resource: str, ns: str, selector, uid, predicate) -> List[Dict[str, Any]]:
        select = getattr(self.store, "select", None)
        if select is not None:      # indexed: cost proportional to the job's own objects
            return select(resource, ns, selector, uid)
        # a store without indexes (HTTP transport, fakes): list and filter
        return [o for o in self.store.list(resource, ns) if predicate(o)]

    def _claim(self, resource: str, job) -> List[Dict[str, Any]]:
        if self.store is None:
            return []
        ns = c.namespace_of(job)
        selector = self.gen_labels(c.name_of(job))
        uid = job.get("metadata", {}).get("uid")
        out = []

        def mine(obj) -> bool:
            # candidates only: carries the job's labels (own or adoptable) or is owned by the job
            # (labels may have drifted: release). Evaluated inside the store before the copy.
            md = obj.get("metadata", {})
            labels = md.get("labels") or {}
            if all(labels.get(k) == v for k, v in selector.items()):
                return True
            return any(r.get("uid") == uid for r in md.get("ownerReferences") or [])
        for obj in self._list_candidates(resource, ns, selector, uid, mine):
            labels = obj.get("metadata", {}).get("labels") or {}
            matches = all(labels.get(k) == v for k, v in selector.items())
            ref = c.get_controller_of(obj)
            if ref is not None:
                if ref.get("uid") != uid:
                    continue          # owned by someone else
                if matches:
                    out.append(obj)
                elif not job.get("metadata", {}).get("deletionTimestamp"):
                    # release: labels drifted away from the selector
                    refs = [r for r in obj["metadata"].get("ownerReferences", [])
                            if r.get("uid") != uid]
                    try:
                        self.store.patch(resource, ns, c.name_of(obj),
                                         {"metadata": {"ownerReferences": refs or None}})
                    except ApiError:
                        pass
                continue
            if not matches or obj.get("metadata", {}).get("deletionTimestamp"):
                continue
            if job.get("metadata", {}).get("deletionTimestamp"):
                continue              # a job being deleted must not adopt (RecheckDeletionTimestamp)
            try:
                adopted = self.store.patch(resource, ns, c.name_of(obj), {
                    "metadata": {"ownerReferences": [self.gen_owner_reference(job)]}})
                out.append(adopted)
            except ApiError:
                pass
        return out

    @staticmethod
    def filter_for_replica_type(objs, rt: str):
        return [o for o in objs
                if (o.get("metadata", {}).get("labels") or {}).get(c.LABEL_REPLICA_TYPE) == rt]

    @staticmethod
    def get_slices(objs, replicas: int) -> List[List[Dict[str, Any]]]:
        """Bucket by replica-index into max(replicas, maxIndex+1) slots (GetPodSlices)."""
        size = replicas
        parsed = []
        for o in objs:
            raw = (o.get("metadata", {}).get("labels") or {}).get(c.LABEL_REPLICA_INDEX)
            try:
                idx = int(raw)
            except (TypeError, ValueError):
                log.warning("object %s has no valid replica-index label", c.name_of(o))
                continue
            if idx < 0:
                continue
            parsed.append((idx, o))
            size = max(size, idx + 1)
        slices: List[List[Dict[str, Any]]] = [[] for _ in range(size)]
        for idx, o in parsed:
            slices[idx].append(o)
        return slices

    # ------------------------------------------------------------------ writes
    def update_job_status_in_api_server(self, job, status) -> None:
        job["status"] = c.deepcopy(status)
        if self.store is None:
            return
        latest = self.store.get(self.info.plural, c.namespace_of(job), c.name_of(job))
        latest["status"] = c.deepcopy(status)
        self.store.update_status(self.info.plural, latest)

    def delete_job(self, job) -> None:
        """tfjob_controller.go:333-349."""
        try:
            if self.store is not None:
                self.store.delete(self.info.plural, c.namespace_of(job), c.name_of(job))
        except ApiError as e:
            if e.code != 404:
                self.recorder.eventf(job, "Warning", REASON_FAILED_DELETE_JOB,
                                     "Error deleting: %s", e.message)
                raise
        self.recorder.eventf(job, "Normal", REASON_SUCCESSFUL_DELETE_JOB, "Deleted job: %s",
                             c.name_of(job))
        if self.metrics is not None:
            self.metrics.jobs_deleted.labels(c.namespace_of(job)).inc()

    # ------------------------------------------------------------------ pods
    def reconcile_pods(self, job, status, pods, rtype: str, spec, replicas) -> None:
        """common.ReconcilePods: slot logic, scale up/down, ExitCode restarts."""
        rt = rtype.lower()
        pods = self.filter_for_replica_type(pods, rt)
        num_replicas = int(spec.get("replicas", 1))
        c.init_replica_status(status, rtype)
        for index, pod_slice in enumerate(self.get_slices(pods, num_replicas)):
            if len(pod_slice) > 1:
                log.warning("We have too many pods for %s %d", rt, index)
            elif len(pod_slice) == 0:
                log.debug("Need to create new pod: %s-%d", rt, index)
                master = self.is_master_role(replicas, rtype, index)
                self.create_new_pod(job, rt, str(index), spec, master, replicas)
            else:
                pod = pod_slice[0]
                if index < 0 or index >= num_replicas:
                    self.pod_control.delete_pod(c.namespace_of(pod), c.name_of(pod), job)
                exit_code = self._record_exit_code(job, pod)
                if spec.get("restartPolicy") == c.RESTART_EXIT_CODE:
                    if (pod.get("status") or {}).get("phase") == c.POD_FAILED and \
                            c.is_retryable_exit_code(exit_code):
                        log.info("Need to restart the pod: %s.%s", c.namespace_of(pod), c.name_of(pod))
                        self.pod_control.delete_pod(c.namespace_of(pod), c.name_of(pod), job)
                        self.on_pod_restart(job, status, rtype)
                c.count_pod_phase(status, rtype, pod)

    def on_pod_restart(self, job, status, rtype: str) -> None:
        """Hook: TFJob appends the Restarting condition here (pod.go:147-157)."""

    def _record_exit_code(self, job, pod) -> int:
        exit_code = 0xBEEF  # magic number, as in the reference
        for cs in (pod.get("status") or {}).get("containerStatuses") or []:
            term = (cs.get("state") or {}).get("terminated")
            if cs.get("name") == self.get_default_container_name() and term is not None:
                exit_code = int(term.get("exitCode", 0))
                self.recorder.eventf(job, "Normal", REASON_EXITED_WITH_CODE,
                                     "Pod: %s.%s exited with code %s", c.namespace_of(pod),
                                     c.name_of(pod), exit_code)
        return exit_code

    def create_new_pod(self, job, rt: str, index: str, spec, master_role: bool, replicas) -> None:
        job_key = c.key_of(job)
        ekey = exp.gen_expectation_pods_key(job_key, rt)
        self.expectations.expect_creations(ekey, 1)
        ref = self.gen_owner_reference(job)
        labels = self.gen_labels(c.name_of(job))
        labels[c.LABEL_REPLICA_TYPE] = rt
        labels[c.LABEL_REPLICA_INDEX] = index
        if master_role:
            labels[c.LABEL_JOB_ROLE] = "master"
        template = c.deepcopy(spec.get("template") or {})
        tmd = template.setdefault("metadata", {})
        tmd["name"] = c.gen_general_name(c.name_of(job), rt, index)
        tlabels = tmd.setdefault("labels", {}) or {}
        tlabels.update(labels)
        tmd["labels"] = tlabels
        pod_spec = template.setdefault("spec", {})
        self.set_cluster_spec(job, template, rt, index)
        if pod_spec.get("restartPolicy"):
            msg = "Restart policy in pod template will be overwritten by restart policy in replica spec"
            log.warning(msg)
            self.recorder.event(job, "Warning", REASON_POD_TEMPLATE_RESTART_POLICY, msg)
        set_restart_policy(template, spec)
        if self.config.enable_gang_scheduling:
            if is_gang_scheduler_set(replicas, self.config.gang_scheduler_name):
                msg = ("Another scheduler is specified when gang-scheduling is enabled and it "
                       "will not be overwritten")
                log.warning(msg)
                self.recorder.event(job, "Warning", REASON_POD_TEMPLATE_SCHEDULER_NAME, msg)
            else:
                pod_spec["schedulerName"] = self.config.gang_scheduler_name
            ann = tmd.setdefault("annotations", {}) or {}
            ann[c.ANNOTATION_POD_GROUP] = c.name_of(job)
            ann[c.ANNOTATION_VOLCANO_TASK_SPEC] = rt
            tmd["annotations"] = ann
        try:
            self.pod_control.create_pods_with_controller_ref(c.namespace_of(job), template, job, ref)
        except TimeoutError:
            # created but initialisation timed out: the informer will observe it later
            return
        except Exception:
            # the watch will never see this pod: roll the expectation back
            self.expectations.creation_observed(ekey)
            raise

    # ------------------------------------------------------------------ services
    def reconcile_services(self, job, services, rtype: str, spec) -> None:
        rt = rtype.lower()
        replicas = int(spec.get("replicas", 1))
        services = self.filter_for_replica_type(services, rt)
        for index, sl in enumerate(self.get_slices(services, replicas)):
            if len(sl) > 1:
                log.warning("We have too many services for %s %d", rt, index)
            elif len(sl) == 0:
                self.create_new_service(job, rtype, spec, str(index))
            else:
                svc = sl[0]
                if index < 0 or index >= replicas:
                    self.service_control.delete_service(c.namespace_of(svc), c.name_of(svc), job)

    def get_ports_from_job(self, spec) -> Dict[str, int]:
        for ctr in c.containers_of(spec):
            if ctr.get("name") == self.get_default_container_name():
                return {p["name"]: int(p["containerPort"]) for p in ctr.get("ports") or []
                        if p.get("name")}
        return {}

    def create_new_service(self, job, rtype: str, spec, index: str) -> None:
        rt = rtype.lower()
        ekey = exp.gen_expectation_services_key(c.key_of(job), rt)
        self.expectations.expect_creations(ekey, 1)
        labels = self.gen_labels(c.name_of(job))
        labels[c.LABEL_REPLICA_TYPE] = rt
        labels[c.LABEL_REPLICA_INDEX] = index
        ports = self.get_ports_from_job(spec)
        service = {
            "apiVersion": "v1", "kind": "Service",
            "metadata": {"name": c.gen_general_name(c.name_of(job), rt, index), "labels": dict(labels)},
            "spec": {"clusterIP": "None", "selector": dict(labels),
                     "ports": [{"name": n, "port": p} for n, p in ports.items()]},
        }
        try:
            self.service_control.create_services_with_controller_ref(
                c.namespace_of(job), service, job, self.gen_owner_reference(job))
        except TimeoutError:
            return
        except Exception:
            self.expectations.creation_observed(ekey)
            raise

    # ------------------------------------------------------------------ policies
    def delete_pods_and_services(self, run_policy, job, pods) -> None:
        """None: nothing; Running: Running+Pending pods; All: every pod. The same-named
        Service goes with each deleted pod (job_test.go:212-289)."""
        if not pods:
            return
        policy = run_policy.get("cleanPodPolicy")
        if policy == c.CLEAN_NONE:
            return
        for pod in pods:
            phase = (pod.get("status") or {}).get("phase")
            if policy == c.CLEAN_RUNNING and phase not in (c.POD_RUNNING, c.POD_PENDING):
                continue
            self.pod_control.delete_pod(c.namespace_of(pod), c.name_of(pod), job)
            self.service_control.delete_service(c.namespace_of(pod), c.name_of(pod), job)

    def cleanup_job(self, run_policy, status, job) -> None:
        """TTL after finished: delete now, or re-queue for the remaining time."""
        ttl = run_policy.get("ttlSecondsAfterFinished")
        if ttl is None:
            return
        done = status.get("completionTime")
        if not done:
            log.warning("job %s finished without a completionTime; cannot apply TTL", c.key_of(job))
            return
        expire = c.parse_time(done) + float(ttl)
        now = self._clock()
        if now >= expire:
            self.delete_job(job)
            return
        self.workqueue.add_after(c.key_of(job), expire - now)

    def past_active_deadline(self, run_policy, status) -> bool:
        ads = run_policy.get("activeDeadlineSeconds")
        start = status.get("startTime")
        if ads is None or not start:
            return False
        return self._clock() - c.parse_time(start) >= float(ads)

    def past_backoff_limit(self, run_policy, replicas, pods) -> bool:
        """Sum of restartCount over init+app containers of Running pods of replica types
        whose restartPolicy is OnFailure/Always (job_test.go:714-737)."""
        limit = run_policy.get("backoffLimit")
        if limit is None:
            return False
        total = 0
        for rtype, spec in replicas.items():
            if spec.get("restartPolicy") not in (c.RESTART_ON_FAILURE, c.RESTART_ALWAYS):
                continue
            for pod in self.filter_for_replica_type(pods, rtype.lower()):
                st = pod.get("status") or {}
                if st.get("phase") != c.POD_RUNNING:
                    continue
                for cs in (st.get("initContainerStatuses") or []) + (st.get("containerStatuses") or []):
                    total += int(cs.get("restartCount", 0))
        if int(limit) == 0:
            return total > 0
        return total >= int(limit)

    # ------------------------------------------------------------------ gang
    def sync_pod_group(self, job, replicas, run_policy) -> Optional[Dict[str, Any]]:
        if self.store is None:
            return None
        ns, name = c.namespace_of(job), c.name_of(job)
        try:
            return self.store.get("podgroups", ns, name)
        except ApiError:
            pass
        sp = run_policy.get("schedulingPolicy") or {}
        min_member = sp.get("minAvailable")
        if min_member is None:
            min_member = c.total_replicas(replicas)
        pg = {
            "apiVersion": "scheduling.volcano.sh/v1beta1", "kind": "PodGroup",
            "metadata": {"name": name, "namespace": ns,
                         "ownerReferences": [self.gen_owner_reference(job)]},
            "spec": {"minMember": int(min_member)},
        }
        if sp.get("queue"):
            pg["spec"]["queue"] = sp["queue"]
        if sp.get("priorityClass"):
            pg["spec"]["priorityClassName"] = sp["priorityClass"]
        if sp.get("minResources"):
            pg["spec"]["minResources"] = c.deepcopy(sp["minResources"])
        return self.store.create("podgroups", pg, ns)

    def delete_pod_group(self, job) -> None:
        if self.store is None:
            return
        try:
            self.store.delete("podgroups", c.namespace_of(job), c.name_of(job))
        except ApiError as e:
            if e.code != 404:
                raise

    # ------------------------------------------------------------------ the state machine
    def reconcile_jobs(self, job, replicas, status, run_policy) -> None:
        job_key = c.key_of(job)
        log.debug("Reconciling for job %s", job_key)
        pods = self.get_pods_for_job(job)
        services = self.get_services_for_job(job)
        old_status = c.deepcopy(status)

        if c.is_finished(status):
            self.delete_pods_and_services(run_policy, job, pods)
            self.cleanup_job(run_policy, status, job)
            if self.config.enable_gang_scheduling:
                self.recorder.event(job, "Normal", "JobTerminated",
                                    "Job has been terminated. Deleting PodGroup")
                self.delete_pod_group(job)
            if c.is_succeeded(status):
                for rs in (status.get("replicaStatuses") or {}).values():
                    if rs.get("active"):
                        rs["succeeded"] = rs.get("succeeded", 0) + rs["active"]
                        rs.pop("active")
            if status != old_status:
                self.update_job_status_in_api_server(job, status)
            return

        # limits
        prev_failed = sum(int(rs.get("failed", 0)) for rs in (status.get("replicaStatuses") or {}).values())
        active = sum(1 for p in pods if (p.get("status") or {}).get("phase") in (c.POD_RUNNING, c.POD_PENDING))
        failed = sum(1 for p in pods if (p.get("status") or {}).get("phase") == c.POD_FAILED)
        total = c.total_replicas(replicas)
        failure_message = ""
        job_exceeds_limit = False
        exceeds_backoff = False
        past_backoff = False
        if run_policy.get("backoffLimit") is not None:
            has_new_failure = failed > prev_failed
            exceeds_backoff = (has_new_failure and active != total and
                               self.workqueue.num_requeues(job_key) + 1 > int(run_policy["backoffLimit"]))
            past_backoff = self.past_backoff_limit(run_policy, replicas, pods)
        if exceeds_backoff or past_backoff:
            job_exceeds_limit = True
            failure_message = (f"Job {c.name_of(job)} has failed because it has reached the "
                               f"specified backoff limit")
        elif self.past_active_deadline(run_policy, status):
            job_exceeds_limit = True
            failure_message = (f"Job {c.name_of(job)} has failed because it was active longer "
                               f"than specified deadline")

        if job_exceeds_limit:
            if status.get("completionTime") is None:
                status["completionTime"] = c.now_rfc3339()
            # a job that ran out of budget takes everything down regardless of cleanPodPolicy
            for pod in pods:
                self.pod_control.delete_pod(c.namespace_of(pod), c.name_of(pod), job)
                self.service_control.delete_service(c.namespace_of(pod), c.name_of(pod), job)
            self.cleanup_job(run_policy, status, job)
            if self.config.enable_gang_scheduling:
                self.recorder.event(job, "Normal", "JobTerminated",
                                    "Job has been terminated. Deleting PodGroup")
                self.delete_pod_group(job)
            self.recorder.event(job, "Normal", c.REASON_JOB_FAILED, failure_message)
            c.update_job_conditions(status, c.JOB_FAILED, c.REASON_JOB_FAILED, failure_message)
            self.on_job_failed(job)
            self.update_job_status_in_api_server(job, status)
            return

        if self.config.enable_gang_scheduling:
            self.sync_pod_group(job, replicas, run_policy)

        for rtype, spec in replicas.items():
            if spec is None:
                continue
            self.reconcile_pods(job, status, pods, rtype, spec, replicas)
            if self.services_for(rtype):
                self.reconcile_services(job, services, rtype, spec)

        self.update_job_status(job, replicas, status)
        if status != old_status:
            status["lastReconcileTime"] = c.now_rfc3339()
            self.update_job_status_in_api_server(job, status)

    def on_job_failed(self, job) -> None:
        """Hook for metrics."""


def set_restart_policy(pod_template, spec) -> None:
    """ExitCode is not a pod restart policy: it maps to Never and the operator does the
    restarting (pod.go:321-328)."""
    pol = spec.get("restartPolicy")
    pod_template.setdefault("spec", {})["restartPolicy"] = \
        c.RESTART_NEVER if pol == c.RESTART_EXIT_CODE else pol


def is_gang_scheduler_set(replicas, scheduler_name: str) -> bool:
    """pkg/common/util/scheduler.go:9-21."""
    scheduler_name = scheduler_name or c.GANG_SCHEDULER_NAME
    for spec in replicas.values():
        sn = (((spec or {}).get("template") or {}).get("spec") or {}).get("schedulerName")
        if sn and sn == scheduler_name:
            return True
    return False
