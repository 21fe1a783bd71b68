"""Job / pod fixtures for the conformance suite — the Python counterparts of
pkg/common/util/v1/testutil/{tfjob,pod,service}.go in the reference."""
import copy

from tf_operator_b200.api import common as c
from tf_operator_b200.engine.control import FakePodControl, FakeServiceControl, Recorder
from tf_operator_b200.engine.job_controller import JobControllerConfig
from tf_operator_b200.store.memory import ObjectStore

TEST_JOB_NAME = "test-tfjob"
TEST_IMAGE = "test-image-for-kubeflow-tf-operator:latest"


def replica_template(container="tensorflow", port_name="tfjob-port", port=2222):
    return {"spec": {"containers": [{"name": container, "image": TEST_IMAGE,
                                     "args": ["Fake", "Fake"],
                                     "ports": [{"name": port_name, "containerPort": port}]}]}}


def new_tfjob(worker, ps, chief=0, master=0, evaluator=0, namespace="default", name=TEST_JOB_NAME,
              **spec_extra):
    specs = {}
    for rt, n in (("Worker", worker), ("PS", ps), ("Chief", chief), ("Master", master),
                  ("Evaluator", evaluator)):
        if n > 0:
            specs[rt] = {"replicas": n, "template": replica_template()}
    job = {"apiVersion": "kubeflow.org/v1", "kind": "TFJob",
           "metadata": {"name": name, "namespace": namespace, "uid": "uid-" + name},
           "spec": {"tfReplicaSpecs": specs}}
    rp = {}
    for k in ("cleanPodPolicy", "ttlSecondsAfterFinished", "activeDeadlineSeconds", "backoffLimit"):
        if k in spec_extra:
            rp[k] = spec_extra.pop(k)
    if rp:
        job["spec"]["runPolicy"] = rp
    job["spec"].update(spec_extra)
    return job


def new_pod(job, typ, index, phase=c.POD_PENDING, exit_code=None, restart_count=None,
            container="tensorflow"):
    pod = {"apiVersion": "v1", "kind": "Pod",
           "metadata": {"name": f"{c.name_of(job)}-{typ}-{index}", "namespace": c.namespace_of(job),
                        "labels": {**c.gen_labels(c.name_of(job)), c.LABEL_REPLICA_TYPE: typ,
                                   c.LABEL_REPLICA_INDEX: str(index)},
                        "ownerReferences": [c.gen_owner_reference(job)]},
           "spec": {}, "status": {"phase": phase}}
    if exit_code is not None:
        pod["status"]["containerStatuses"] = [{"name": container,
                                               "state": {"terminated": {"exitCode": exit_code}}}]
    if restart_count is not None:
        pod["status"].setdefault("containerStatuses", [{"name": container}])
        pod["status"]["containerStatuses"][0]["restartCount"] = restart_count
    return pod


def set_pods_statuses(store, job, typ, pending, active, succeeded, failed, restart_counts=None):
    """Same index order as testutil.SetPodsStatuses: pending, running, succeeded, failed."""
    idx = 0
    for phase, n in ((c.POD_PENDING, pending), (c.POD_RUNNING, active),
                     (c.POD_SUCCEEDED, succeeded), (c.POD_FAILED, failed)):
        for i in range(n):
            rc = restart_counts[i] if (restart_counts and phase == c.POD_RUNNING) else None
            pod = new_pod(job, typ, idx, phase, restart_count=rc)
            created = store.create("pods", pod)
            created["status"] = pod["status"]
            store.update_status("pods", created)
            idx += 1


def set_services(store, job, typ, n):
    for i in range(n):
        store.create("services", {
            "metadata": {"name": f"{c.name_of(job)}-{typ}-{i}", "namespace": c.namespace_of(job),
                         "labels": {**c.gen_labels(c.name_of(job)), c.LABEL_REPLICA_TYPE: typ,
                                    c.LABEL_REPLICA_INDEX: str(i)},
                         "ownerReferences": [c.gen_owner_reference(job)]},
            "spec": {"clusterIP": "None"}})


def make_controller(cls, job=None, gang=False):
    """Reconciler wired like the reference's unit tests: reads from a hand-filled store,
    writes go to Fake controls that only record."""
    store = ObjectStore()
    ctrl = cls(store, pod_control=FakePodControl(), service_control=FakeServiceControl(),
               recorder=Recorder(None, "test"),
               config=JobControllerConfig(enable_gang_scheduling=gang))
    if job is not None:
        created = store.create(ctrl.info.plural, copy.deepcopy(job))
        job["metadata"]["uid"] = created["metadata"]["uid"]
        job["metadata"]["resourceVersion"] = created["metadata"]["resourceVersion"]
    return ctrl, store
