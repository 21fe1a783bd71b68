"""API server semantics (store/), REST façade, SDK surface, rendezvous store, name mapping,
topology placement — the CPU-side plumbing under the operator."""
import io
import json
import threading
import time
import urllib.request

import pytest

from tf_operator_b200.api import common as c
from tf_operator_b200.cluster import LocalCluster
from tf_operator_b200.launcher import netmap
from tf_operator_b200.launcher.rendezvous import RendezvousClient, RendezvousServer
from tf_operator_b200.launcher.topology import GpuInfo, Topology, synthetic
from tf_operator_b200.sdk import (PyTorchJobClient, TFJobClient, V1Container, V1ObjectMeta, V1PodSpec,
                                  V1PodTemplateSpec, V1PyTorchJob, V1PyTorchJobSpec, V1ReplicaSpec,
                                  V1TFJob, V1TFJobSpec, utils, watch)
from tf_operator_b200.sdk.models import to_wire
from tf_operator_b200.sdk.transport import HttpTransport
from tf_operator_b200.store.http import ApiServer
from tf_operator_b200.store.memory import ApiError, ObjectStore, labels_match, parse_label_selector


def _pod(name, labels=None, owner=None):
    md = {"name": name, "namespace": "default", "labels": labels or {}}
    if owner:
        md["ownerReferences"] = [owner]
    return {"metadata": md, "spec": {}}


# ---- store ------------------------------------------------------------------------------------
def test_store_crud_versions_and_conflict():
    s = ObjectStore()
    a = s.create("pods", _pod("a"))
    assert a["metadata"]["uid"] and a["metadata"]["resourceVersion"] == "1" and a["status"]["phase"] == "Pending"
    with pytest.raises(ApiError) as e:
        s.create("pods", _pod("a"))
    assert e.value.code == 409 and e.value.reason == "AlreadyExists"
    a["spec"]["x"] = 1
    b = s.update("pods", a)
    assert int(b["metadata"]["resourceVersion"]) > 1 and b["metadata"]["generation"] == 2
    with pytest.raises(ApiError) as e:
        s.update("pods", a)                       # stale resourceVersion
    assert e.value.reason == "Conflict"
    with pytest.raises(ApiError) as e:
        s.get("pods", "default", "nope")
    assert e.value.code == 404


def test_status_subresource_is_separate_from_spec():
    s = ObjectStore()
    p = s.create("pods", _pod("p"))
    p["status"] = {"phase": "Running"}
    p["spec"]["changed"] = True
    out = s.update_status("pods", p)
    assert out["status"]["phase"] == "Running" and "changed" not in out["spec"]
    out["spec"]["changed"] = True
    out["status"] = {"phase": "Failed"}
    out2 = s.update("pods", out)                  # spec write keeps the stored status
    assert out2["spec"]["changed"] and out2["status"]["phase"] == "Running"


def test_selectors_watch_and_cascade_delete():
    s = ObjectStore()
    assert labels_match({"a": "1", "b": "2"}, parse_label_selector("a=1,b!=3,a"))
    assert not labels_match({"a": "1"}, parse_label_selector("!a"))
    job = s.create("tfjobs", {"kind": "TFJob", "metadata": {"name": "j"}, "spec": {"tfReplicaSpecs": {}}})
    ref = c.gen_owner_reference(job)
    w = s.watch("pods", "default", "job-name=j")
    s.create("pods", _pod("p0", {"job-name": "j"}, ref))
    s.create("pods", _pod("other", {"job-name": "k"}))
    s.create("services", _pod("p0", {"job-name": "j"}, ref))
    assert [c.name_of(p) for p in s.list("pods", "default", "job-name=j")] == ["p0"]
    assert s.list("pods", "default", field_selector="metadata.name=other")[0]["metadata"]["name"] == "other"
    ev = w.get(timeout=1)
    assert ev[0] == "ADDED" and c.name_of(ev[1]) == "p0"
    s.delete("tfjobs", "default", "j")            # ownerReference GC
    assert [c.name_of(p) for p in s.list("pods")] == ["other"] and s.list("services") == []
    ev = w.get(timeout=1)
    assert ev[0] == "DELETED"
    w.stop()


def test_admission_of_job_kinds():
    s = ObjectStore()
    with pytest.raises(ApiError) as e:
        s.create("tfjobs", {"metadata": {"name": "x"}, "spec": {}})
    assert e.value.code == 422 and "spec.tfReplicaSpecs: Required value" in e.value.message
    with pytest.raises(ApiError) as e:
        s.create("mxjobs", {"metadata": {"name": "x"}, "spec": {"mxReplicaSpecs": {}}})
    assert "spec.jobMode: Required value" in e.value.message
    with pytest.raises(ApiError) as e:
        s.create("pytorchjobs", {"metadata": {"name": "x"}, "spec": {"pytorchReplicaSpecs": {"Master": {"replicas": "2"}}}})
    assert "must be an integer" in e.value.message
    ok = s.create("tfjobs", {"metadata": {"name": "flat"}, "spec": {"cleanPodPolicy": "All", "tfReplicaSpecs": {}}})
    assert ok["spec"]["runPolicy"]["cleanPodPolicy"] == "All" and ok["kind"] == "TFJob"


def test_merge_patch():
    s = ObjectStore()
    s.create("pods", _pod("p", {"a": "1", "b": "2"}))
    out = s.patch("pods", "default", "p", {"metadata": {"labels": {"a": None, "c": "3"}}})
    assert out["metadata"]["labels"] == {"b": "2", "c": "3"}


# ---- REST façade -------------------------------------------------------------------------------
def test_http_facade_crud_watch_and_errors():
    s = ObjectStore()
    api = ApiServer(s).start()
    try:
        t = HttpTransport(api.url)
        job = {"apiVersion": "kubeflow.org/v1", "kind": "PyTorchJob", "metadata": {"name": "h"},
               "spec": {"pytorchReplicaSpecs": {"Master": {"replicas": 1}}}}
        assert t.create("pytorchjobs", "default", job)["metadata"]["name"] == "h"
        assert t.get("pytorchjobs", "default", "h")["kind"] == "PyTorchJob"
        assert len(t.list("pytorchjobs", "default")) == 1
        t.patch("pytorchjobs", "default", "h", {"metadata": {"labels": {"x": "y"}}})
        assert t.list("pytorchjobs", "default", "x=y")[0]["metadata"]["labels"] == {"x": "y"}
        raw = json.loads(urllib.request.urlopen(api.url + "/apis/kubeflow.org/v1/namespaces/default/pytorchjobs").read())
        assert raw["kind"] == "PyTorchJobList" and raw["items"]
        assert urllib.request.urlopen(api.url + "/healthz").read() == b"ok"
        got = []
        th = threading.Thread(target=lambda: got.extend(
            e for e in t.watch("pytorchjobs", "default", 2)), daemon=True)
        th.start()
        time.sleep(0.3)
        t.delete("pytorchjobs", "default", "h")
        th.join(4)
        assert [e[0] for e in got][:1] == ["ADDED"] and "DELETED" in [e[0] for e in got]
        with pytest.raises(ApiError) as e:
            t.get("pytorchjobs", "default", "h")
        assert e.value.code == 404
    finally:
        api.stop()


# ---- SDK ------------------------------------------------------------------------------------------
def test_models_wire_format_and_equality():
    ctr = V1Container(name="tensorflow", image="img", args=["--x"], working_dir="/w")
    spec = V1TFJobSpec(clean_pod_policy="All", ttl_seconds_after_finished=3,
                       tf_replica_specs={"Worker": V1ReplicaSpec(replicas=2, restart_policy="OnFailure",
                                                                 template=V1PodTemplateSpec(spec=V1PodSpec(containers=[ctr])))})
    job = V1TFJob(api_version="kubeflow.org/v1", kind="TFJob", metadata=V1ObjectMeta(name="j", namespace="ns"), spec=spec)
    wire = to_wire(job)
    assert wire["apiVersion"] == "kubeflow.org/v1" and wire["metadata"] == {"name": "j", "namespace": "ns"}
    assert wire["spec"]["cleanPodPolicy"] == "All" and wire["spec"]["ttlSecondsAfterFinished"] == 3
    w = wire["spec"]["tfReplicaSpecs"]["Worker"]
    assert w["restartPolicy"] == "OnFailure" and w["template"]["spec"]["containers"][0]["workingDir"] == "/w"
    assert job == V1TFJob(api_version="kubeflow.org/v1", kind="TFJob", metadata=V1ObjectMeta(name="j", namespace="ns"), spec=spec)
    assert job != V1TFJob(kind="TFJob")
    assert "tf_replica_specs" in job.spec.to_dict() and "Worker" in job.to_str()
    assert V1TFJobSpec.attribute_map["tf_replica_specs"] == "tfReplicaSpecs"
    assert utils.set_tfjob_namespace(job) == "ns" and utils.set_job_namespace({"metadata": {}}) == "default"
    assert utils.to_selector(utils.get_labels("j", master=True, replica_type="Worker", replica_index=0)) == \
        "group-name=kubeflow.org,job-name=j,job-role=master,replica-type=worker,replica-index=0"
    pj = V1PyTorchJob(kind="PyTorchJob", metadata=V1ObjectMeta(name="p"),
                      spec=V1PyTorchJobSpec(pytorch_replica_specs={"Master": V1ReplicaSpec(replicas=1)}))
    assert to_wire(pj)["spec"]["pytorchReplicaSpecs"]["Master"]["replicas"] == 1


def test_sdk_end_to_end_like_the_reference_e2e():
    """sdk/python/test/test_e2e.py:30-72: create -> wait_for_job -> is_job_succeeded -> get_logs -> delete,
    plus watch and the timeout error."""
    with LocalCluster(topology=synthetic(0)) as cl:
        cli = TFJobClient(cluster=cl)
        ctr = V1Container(name="tensorflow", image="x", command=["python", "-c", "print('mnist done')"])
        job = V1TFJob(api_version="kubeflow.org/v1", kind="TFJob", metadata=V1ObjectMeta(name="mnist", namespace="default"),
                      spec=V1TFJobSpec(clean_pod_policy="None", tf_replica_specs={"Worker": V1ReplicaSpec(
                          replicas=1, restart_policy="Never", template=V1PodTemplateSpec(spec=V1PodSpec(containers=[ctr])))}))
        cli.create(job)
        seen = []
        cli.wait_for_job("mnist", polling_interval=0.05, timeout_seconds=30, status_callback=seen.append)
        assert seen and cli.is_job_succeeded("mnist") and not cli.is_job_running("mnist")
        assert cli.get_job_status("mnist") == "Succeeded"
        assert "mnist done" in cli.get_logs("mnist", master=False)["mnist-worker-0"]
        assert cli.get()["items"][0]["metadata"]["name"] == "mnist"
        buf = io.StringIO()
        watch(cli, name="mnist", namespace="default", timeout_seconds=5, out=buf)
        assert "NAME" in buf.getvalue() and "Succeeded" in buf.getvalue()
        cli.patch("mnist", {"metadata": {"labels": {"team": "a"}}})
        assert cli.get("mnist")["metadata"]["labels"]["team"] == "a"
        with pytest.raises(RuntimeError, match="Timeout waiting"):
            cli.wait_for_condition("mnist", ["NeverHappens"], timeout_seconds=0.2, polling_interval=0.1)
        cli.delete("mnist")
        with pytest.raises(RuntimeError):
            cli.get("mnist")
        with pytest.raises(RuntimeError, match="Not found Pods"):
            cli.get_logs("mnist")
        assert PyTorchJobClient(cluster=cl).get()["items"] == []


# ---- launcher plumbing ------------------------------------------------------------------------------
def test_netmap_is_deterministic_and_parses_service_names():
    ip = netmap.pod_ip("default", "job-worker-0")
    assert ip == netmap.pod_ip("default", "job-worker-0") and ip.startswith("127.") and not ip.startswith("127.0.0.")
    assert netmap.resolve("job-worker-0.default.svc") == ip
    assert netmap.resolve("job-worker-0.default.svc.cluster.local", "other") == ip
    assert netmap.resolve("job-worker-0", "default") == ip
    assert netmap.resolve("localhost") == "localhost" and netmap.resolve("10.1.2.3") == "10.1.2.3"
    assert netmap.resolve_endpoint("job-worker-0.default.svc:2222") == ip + ":2222"
    assert netmap.pod_ip("ns2", "job-worker-0") != ip


def test_rendezvous_store_all_gather_barrier_and_timeout():
    srv = RendezvousServer().start()
    try:
        out = {}

        def rank(r):
            cli = RendezvousClient(srv.addr)
            out[r] = cli.all_gather("job/ipc", r, 3, bytes([r]) * 4)
            cli.barrier("job/ready", 3)
            cli.barrier("job/ready", 3)          # reusable
            cli.close()
        ths = [threading.Thread(target=rank, args=(r,)) for r in range(3)]
        [t.start() for t in ths]
        [t.join(10) for t in ths]
        assert all(out[r] == [b"\x00" * 4, b"\x01" * 4, b"\x02" * 4] for r in range(3))
        cli = RendezvousClient(srv.addr)
        assert cli.add("ctr", 5) == 5 and cli.add("ctr") == 6
        with pytest.raises(TimeoutError):
            cli.get("missing", timeout=0.1)
        assert cli.delete_prefix("job/") >= 3
        cli.close()
    finally:
        srv.stop()


def test_topology_placement():
    t = synthetic(8)
    assert t.pick(3, [5, 1, 7, 2]) == [1, 2, 5] and t.pick(9, list(range(8))) is None and t.pick(0, []) == []
    # no switch: GPUs 0-1 and 2-3 are NVLink pairs; a 2-GPU job must get a pair
    pairs = Topology(gpus=[GpuInfo(0, nvlink_peers={1: 4}), GpuInfo(1, nvlink_peers={0: 4}),
                           GpuInfo(2, nvlink_peers={3: 4}), GpuInfo(3, nvlink_peers={2: 4})])
    assert pairs.pick(2, [0, 2, 3]) == [2, 3]
    assert pairs.pick(2, [0, 1, 2]) == [0, 1]


def test_swagger_runtime_standins_roundtrip_models_over_rest():
    """ApiClient / Configuration / rest: raw REST with the SDK models against
    the REST facade - serialise a V1TFJob to camelCase JSON, POST, GET back as a model."""
    from tf_operator_b200.cluster import LocalCluster
    from tf_operator_b200.launcher.topology import synthetic
    from tf_operator_b200.sdk import (ApiClient, ApiException, Configuration, V1Container, V1ObjectMeta,
                                      V1PodSpec, V1PodTemplateSpec, V1ReplicaSpec, V1TFJob, V1TFJobSpec)
    with LocalCluster(topology=synthetic(0), serve_http=True) as cl:
        cfg = Configuration()
        cfg.host = cl.api.url
        Configuration.set_default(cfg)
        api = ApiClient()
        assert api.configuration.host == cl.api.url and "User-Agent" in api.default_headers
        job = V1TFJob(api_version="kubeflow.org/v1", kind="TFJob",
                      metadata=V1ObjectMeta(name="raw", namespace="default"),
                      spec=V1TFJobSpec(tf_replica_specs={"Worker": V1ReplicaSpec(
                          replicas=2, restart_policy="Never", template=V1PodTemplateSpec(spec=V1PodSpec(
                              containers=[V1Container(name="tensorflow", image="x", command=["true"])])))}))
        wire = api.sanitize_for_serialization(job)
        assert wire["spec"]["tfReplicaSpecs"]["Worker"]["restartPolicy"] == "Never"
        path = "/apis/kubeflow.org/v1/namespaces/{namespace}/tfjobs"
        created = api.call_api(path, "POST", path_params={"namespace": "default"}, body=job,
                               response_type="V1TFJob")
        assert isinstance(created, V1TFJob) and created.metadata.name == "raw"
        got, status, _ = api.call_api(path + "/{name}", "GET", path_params={"namespace": "default", "name": "raw"},
                                      response_type="V1TFJob", _return_http_data_only=False)
        assert status == 200 and got.spec.tf_replica_specs["Worker"].replicas == 2
        lst = api.call_api(path, "GET", path_params={"namespace": "default"}, response_type="V1TFJobList")
        assert [j.metadata.name for j in lst.items] == ["raw"]
        import pytest
        with pytest.raises(ApiException) as ei:
            api.call_api(path + "/{name}", "GET", path_params={"namespace": "default", "name": "nope"})
        assert ei.value.status == 404
        assert api.select_header_accept(["application/yaml", "application/json"]) == "application/json"
        assert api.select_header_content_type([]) == "application/json"
        assert "SDK Package Version" in cfg.to_debug_report()
    Configuration.set_default(None)


def test_store_indexes_agree_with_a_full_scan_under_random_mutation():
    """ObjectStore.select (label and owner indexes) returns exactly what a walk over every object
    would, after any mix of creates, label / owner edits, status updates and (cascading) deletes;
    ``list(predicate=...)`` filters before it copies."""
    import random
    from tf_operator_b200.store.memory import ApiError, ObjectStore
    rng = random.Random(7)
    st = ObjectStore()
    owners = [f"uid-{i}" for i in range(4)]
    names = [f"p{i}" for i in range(40)]

    def brute(labels, uid):
        out = []
        for o in st.list("pods", "default"):
            md = o["metadata"]
            lab = md.get("labels") or {}
            if (labels and all(lab.get(k) == v for k, v in labels.items())) or \
                    any(r.get("uid") == uid for r in md.get("ownerReferences") or []):
                out.append(md["name"])
        return sorted(out)

    for step in range(600):
        name = rng.choice(names)
        op = rng.random()
        try:
            if op < 0.35:
                md = {"name": name, "namespace": "default",
                      "labels": {"job-name": f"j{rng.randrange(5)}", "replica-type": rng.choice(["ps", "worker"])}}
                if rng.random() < 0.7:
                    md["ownerReferences"] = [{"uid": rng.choice(owners), "controller": True, "kind": "TFJob",
                                              "name": "x", "apiVersion": "kubeflow.org/v1"}]
                st.create("pods", {"metadata": md, "spec": {}})
            elif op < 0.55:
                st.patch("pods", "default", name, {"metadata": {"labels": {"job-name": f"j{rng.randrange(5)}"}}})
            elif op < 0.65:
                st.patch("pods", "default", name, {"metadata": {"ownerReferences": None}})
            elif op < 0.8:
                p = st.get("pods", "default", name)
                p.setdefault("status", {})["phase"] = "Running"
                st.update_status("pods", p)
            else:
                st.delete("pods", "default", name)
        except ApiError:
            pass
        if step % 10 == 0:
            for j in range(5):
                sel = {"job-name": f"j{j}"}
                uid = rng.choice(owners)
                assert [o["metadata"]["name"] for o in st.select("pods", "default", sel, uid)] == brute(sel, uid)
                assert [o["metadata"]["name"] for o in st.select("pods", "default", sel)] == brute(sel, None)
                two = {"job-name": f"j{j}", "replica-type": "ps"}
                assert [o["metadata"]["name"] for o in st.select("pods", "default", two)] == brute(two, None)
    # the predicate form of list copies only what it keeps, and sees the stored labels
    hits = st.list("pods", "default", predicate=lambda o: (o["metadata"].get("labels") or {}).get("job-name") == "j1")
    assert sorted(o["metadata"]["name"] for o in hits) == brute({"job-name": "j1"}, None)
    # a cascading delete goes through the owner index
    owner = st.create("tfjobs", {"apiVersion": "kubeflow.org/v1", "kind": "TFJob",
                                 "metadata": {"name": "own", "namespace": "default"},
                                 "spec": {"tfReplicaSpecs": {"Worker": {"replicas": 1, "template": {"spec": {
                                     "containers": [{"name": "tensorflow", "image": "i"}]}}}}}})
    for i in range(3):
        st.create("pods", {"metadata": {"name": f"dep{i}", "namespace": "default", "ownerReferences": [
            {"uid": owner["metadata"]["uid"], "controller": True, "kind": "TFJob", "name": "own",
             "apiVersion": "kubeflow.org/v1"}]}, "spec": {}})
    st.delete("tfjobs", "default", "own")
    assert not [p for p in st.list("pods", "default") if p["metadata"]["name"].startswith("dep")]
    assert st.select("pods", "default", {}, owner["metadata"]["uid"]) == []


def test_store_keeps_a_bounded_number_of_events():
    from tf_operator_b200.store.memory import ObjectStore
    st = ObjectStore()
    st.max_events = 50
    pod = st.create("pods", {"metadata": {"name": "p", "namespace": "default"}, "spec": {}})
    for i in range(180):
        st.record_event(pod, "Normal", "Tick", f"event {i}")
    evs = st.list("events", "default")
    assert len(evs) == 50 and {e["message"] for e in evs} == {f"event {i}" for i in range(130, 180)}
    assert len(st.audit) <= st.audit.maxlen
