"""API defaults / validation tables for the four kinds (pkg/apis/*/v1/defaults_test.go,
pkg/apis/*/validation/validation_test.go, pkg/apis/tensorflow/v1/util_test.go)."""
import copy

import pytest
import yaml

from tf_operator_b200 import api
from tf_operator_b200.api import common as c
from tf_operator_b200.api import mxnet as mx
from tf_operator_b200.api import pytorch as pt
from tf_operator_b200.api import tensorflow as tf
from tf_operator_b200.api import xgboost as xgb

IMG = "test-image-for-kubeflow-tf-operator:latest"


def rspec(container, image=IMG, ports=None, replicas=None, policy=None):
    ctr = {"name": container}
    if image is not None:
        ctr["image"] = image
    if ports is not None:
        ctr["ports"] = ports
    out = {"template": {"spec": {"containers": [ctr]}}}
    if replicas is not None:
        out["replicas"] = replicas
    if policy:
        out["restartPolicy"] = policy
    return out


# defaults_test.go:83-147 TestSetTypeNames ---------------------------------------------------
@pytest.mark.parametrize("given,want", [("ps", "PS"), ("WORKER", "Worker"), ("chief", "Chief"),
                                        ("mAsTeR", "Master"), ("evaluator", "Evaluator")])
def test_tf_type_names_normalised(given, want):
    job = {"kind": "TFJob", "spec": {"tfReplicaSpecs": {given: rspec("tensorflow")}}}
    tf.set_defaults(job)
    assert list(job["spec"]["tfReplicaSpecs"]) == [want]


# defaults_test.go:149-283 TestSetDefaultTFJob ---------------------------------------------------
def test_tf_defaults_replicas_policy_port_cleanpod():
    job = {"kind": "TFJob", "spec": {"tfReplicaSpecs": {"Worker": rspec("tensorflow")}}}
    tf.set_defaults(job)
    w = job["spec"]["tfReplicaSpecs"]["Worker"]
    assert w["replicas"] == 1 and w["restartPolicy"] == "Never"
    assert w["template"]["spec"]["containers"][0]["ports"] == [{"name": "tfjob-port", "containerPort": 2222}]
    assert job["spec"]["runPolicy"]["cleanPodPolicy"] == "Running"
    assert job["spec"]["successPolicy"] == ""


def test_tf_default_port_appended_next_to_custom_port_and_kept_when_named():
    custom = [{"name": "customPort", "containerPort": 1234}]
    job = {"kind": "TFJob", "spec": {"tfReplicaSpecs": {"Worker": rspec("tensorflow", ports=custom)}}}
    tf.set_defaults(job)
    assert job["spec"]["tfReplicaSpecs"]["Worker"]["template"]["spec"]["containers"][0]["ports"] == [
        {"name": "customPort", "containerPort": 1234}, {"name": "tfjob-port", "containerPort": 2222}]
    named = [{"name": "tfjob-port", "containerPort": 9999}]
    job = {"kind": "TFJob", "spec": {"tfReplicaSpecs": {"Worker": rspec("tensorflow", ports=named)}}}
    tf.set_defaults(job)
    assert job["spec"]["tfReplicaSpecs"]["Worker"]["template"]["spec"]["containers"][0]["ports"] == named


def test_tf_default_port_goes_to_named_container_else_first():
    spec = {"template": {"spec": {"containers": [{"name": "sidecar", "image": IMG},
                                                 {"name": "tensorflow", "image": IMG}]}}}
    job = {"kind": "TFJob", "spec": {"tfReplicaSpecs": {"Worker": spec}}}
    tf.set_defaults(job)
    ctrs = job["spec"]["tfReplicaSpecs"]["Worker"]["template"]["spec"]["containers"]
    assert "ports" not in ctrs[0] and ctrs[1]["ports"][0]["name"] == "tfjob-port"
    spec = {"template": {"spec": {"containers": [{"name": "a", "image": IMG}, {"name": "b", "image": IMG}]}}}
    job = {"kind": "TFJob", "spec": {"tfReplicaSpecs": {"Worker": spec}}}
    tf.set_defaults(job)
    assert "ports" in job["spec"]["tfReplicaSpecs"]["Worker"]["template"]["spec"]["containers"][0]


def test_explicit_values_survive_defaulting():
    job = {"kind": "TFJob", "spec": {"runPolicy": {"cleanPodPolicy": "All"}, "successPolicy": "AllWorkers",
                                     "tfReplicaSpecs": {"Worker": rspec("tensorflow", replicas=3, policy="ExitCode")}}}
    tf.set_defaults(job)
    assert job["spec"]["runPolicy"]["cleanPodPolicy"] == "All"
    assert job["spec"]["successPolicy"] == "AllWorkers"
    assert job["spec"]["tfReplicaSpecs"]["Worker"]["replicas"] == 3
    assert job["spec"]["tfReplicaSpecs"]["Worker"]["restartPolicy"] == "ExitCode"


def test_flat_pre_runpolicy_spec_is_lifted():
    """examples/v1/mnist_with_summaries/tf_job_mnist.yaml:7 style."""
    job = {"kind": "TFJob", "spec": {"cleanPodPolicy": "None", "ttlSecondsAfterFinished": 5,
                                     "tfReplicaSpecs": {"Worker": rspec("tensorflow")}}}
    tf.set_defaults(job)
    assert job["spec"]["runPolicy"] == {"cleanPodPolicy": "None", "ttlSecondsAfterFinished": 5}
    assert "cleanPodPolicy" not in job["spec"]


# util_test.go:23-48 ---------------------------------------------------------------------------
def test_is_chief_or_master():
    assert tf.is_chief_or_master("Chief") and tf.is_chief_or_master("Master")
    assert not tf.is_chief_or_master("Worker") and not tf.is_chief_or_master("PS")
    assert tf.is_worker("Worker") and tf.is_evaluator("Evaluator")


# per-kind defaults ----------------------------------------------------------------------------
def test_pytorch_defaults_port_only_on_master():
    job = {"kind": "PyTorchJob", "spec": {"pytorchReplicaSpecs": {
        "master": rspec("pytorch"), "WORKER": rspec("pytorch", replicas=2)}}}
    pt.set_defaults(job)
    specs = job["spec"]["pytorchReplicaSpecs"]
    assert set(specs) == {"Master", "Worker"}
    assert specs["Master"]["template"]["spec"]["containers"][0]["ports"] == [
        {"name": "pytorchjob-port", "containerPort": 23456}]
    assert "ports" not in specs["Worker"]["template"]["spec"]["containers"][0]
    assert specs["Master"]["restartPolicy"] == "OnFailure" and specs["Master"]["replicas"] == 1
    assert job["spec"]["runPolicy"]["cleanPodPolicy"] == "None"


def test_mxnet_defaults():
    job = {"kind": "MXJob", "spec": {"jobMode": "MXTrain", "mxReplicaSpecs": {
        "scheduler": rspec("mxnet"), "SERVER": rspec("mxnet"), "worker": rspec("mxnet"),
        "tuner": rspec("mxnet")}}}
    mx.set_defaults(job)
    specs = job["spec"]["mxReplicaSpecs"]
    # only Scheduler/Server/Worker keys are case-normalised (defaults.go:71-75)
    assert set(specs) == {"Scheduler", "Server", "Worker", "tuner"}
    for s in specs.values():
        assert s["restartPolicy"] == "Never" and s["replicas"] == 1
        assert s["template"]["spec"]["containers"][0]["ports"] == [{"name": "mxjob-port", "containerPort": 9091}]
    assert job["spec"]["runPolicy"]["cleanPodPolicy"] == "All"


def test_xgboost_defaults():
    job = {"kind": "XGBoostJob", "spec": {"xgbReplicaSpecs": {"master": rspec("xgboost"),
                                                               "worker": rspec("xgboost", replicas=2)}}}
    xgb.set_defaults(job)
    specs = job["spec"]["xgbReplicaSpecs"]
    assert set(specs) == {"Master", "Worker"}
    for s in specs.values():
        assert s["restartPolicy"] == "Never"
        assert s["template"]["spec"]["containers"][0]["ports"] == [{"name": "xgboostjob-port", "containerPort": 9999}]
    assert job["spec"]["runPolicy"]["cleanPodPolicy"] == "All"


# validation tables ----------------------------------------------------------------------------
@pytest.mark.parametrize("spec,msg", [
    ({}, "TFJobSpec is not valid"),
    ({"tfReplicaSpecs": {"Worker": {"template": {"spec": {"containers": []}}}}},
     "TFJobSpec is not valid: containers definition expected in Worker"),
    ({"tfReplicaSpecs": {"Worker": rspec("tensorflow", image="")}},
     "TFJobSpec is not valid: Image is undefined in the container of Worker"),
    ({"tfReplicaSpecs": {"Worker": rspec("", image=IMG)}},
     "TFJobSpec is not valid: There is no container named tensorflow in Worker"),
    ({"tfReplicaSpecs": {"Chief": rspec("tensorflow"), "Master": rspec("tensorflow")}},
     "TFJobSpec is not valid: more than 1 chief/master found"),
])
def test_tf_validation_rejects(spec, msg):
    with pytest.raises(c.ValidationError) as e:
        tf.validate(spec)
    assert str(e.value) == msg


def test_tf_validation_accepts():
    tf.validate({"tfReplicaSpecs": {"Worker": rspec("tensorflow"), "PS": rspec("tensorflow"),
                                    "Chief": rspec("tensorflow")}})


@pytest.mark.parametrize("spec,msg", [
    ({}, "PyTorchJobSpec is not valid"),
    ({"pytorchReplicaSpecs": {"Master": {"template": {"spec": {"containers": []}}}}},
     "PyTorchJobSpec is not valid: containers definition expected in Master"),
    ({"pytorchReplicaSpecs": {"Master": rspec("pytorch", image="")}},
     "PyTorchJobSpec is not valid: Image is undefined in the container of Master"),
    ({"pytorchReplicaSpecs": {"Master": rspec("", image=IMG)}},
     "PyTorchJobSpec is not valid: There is no container named pytorch in Master"),
    ({"pytorchReplicaSpecs": {"Master": rspec("pytorch", replicas=2)}},
     "PyTorchJobSpec is not valid: There must be only 1 master replica"),
    ({"pytorchReplicaSpecs": {"Worker": rspec("pytorch")}},
     "PyTorchJobSpec is not valid: Master ReplicaSpec must be present"),
    ({"pytorchReplicaSpecs": {"Master": rspec("pytorch"), "PS": rspec("pytorch")}},
     "PyTorchReplicaType is PS but must be one of [Master Worker]"),
])
def test_pytorch_validation_rejects(spec, msg):
    with pytest.raises(c.ValidationError) as e:
        pt.validate(spec)
    assert str(e.value) == msg


@pytest.mark.parametrize("spec,msg", [
    ({}, "XGBoostJobSpec is not valid"),
    ({"xgbReplicaSpecs": {"Master": {"template": {"spec": {"containers": []}}}}},
     "XGBoostJobSpec is not valid: containers definition expected in Master"),
    ({"xgbReplicaSpecs": {"Master": rspec("xgboost", image="")}},
     "XGBoostReplicaType is not valid: Image is undefined in the container of Master"),
    ({"xgbReplicaSpecs": {"Master": rspec("", image=IMG)}},
     "XGBoostReplicaType is not valid: There is no container named xgboost in Master"),
    ({"xgbReplicaSpecs": {"Master": rspec("xgboost", replicas=2)}},
     "XGBoostReplicaType is not valid: There must be only 1 master replica"),
    ({"xgbReplicaSpecs": {"Worker": rspec("xgboost")}},
     "XGBoostReplicaType is not valid: Master ReplicaSpec must be present"),
])
def test_xgboost_validation_rejects(spec, msg):
    with pytest.raises(c.ValidationError) as e:
        xgb.validate(spec)
    assert str(e.value) == msg


@pytest.mark.parametrize("spec,msg", [
    ({}, "MXJobSpec is not valid"),
    ({"mxReplicaSpecs": {"Worker": {"template": {"spec": {"containers": []}}}}}, "MXJobSpec is not valid"),
    ({"mxReplicaSpecs": {"Worker": rspec("mxnet", image="")}}, "MXJobSpec is not valid"),
    ({"mxReplicaSpecs": {"Worker": rspec("", image=IMG)}}, "MXJobSpec is not valid"),
])
def test_mxnet_validation_rejects(spec, msg):
    with pytest.raises(c.ValidationError) as e:
        mx.validate(spec)
    assert str(e.value) == msg


def test_mxnet_validation_accepts_one_scheduler():
    mx.validate({"mxReplicaSpecs": {"Scheduler": rspec("mxnet"), "Server": rspec("mxnet"),
                                    "Worker": rspec("mxnet")}})


# the reference's own example manifests load, default and validate ---------------------------------
@pytest.mark.parametrize("path", [
    "examples/v1/dist-mnist/tf_job_mnist.yaml",
    "examples/v1/mnist_with_summaries/tf_job_mnist.yaml",
    "examples/v1/distribution_strategy/keras-API/multi_worker_tfjob.yaml",
    "examples/v1/distribution_strategy/estimator-API/distributed_tfjob.yaml",
    "examples/mxnet/mxjob_dist_v1.yaml",
    "examples/xgboost/xgboostjob.yaml",
])
def test_reference_example_manifests(path):
    import os
    full = os.path.join("/root/reference", path)
    if not os.path.exists(full):
        pytest.skip("reference tree not mounted")
    for doc in yaml.safe_load_all(open(full)):
        if not doc or doc.get("kind") not in api.KINDS:
            continue
        api.set_defaults(doc)
        api.validate(doc)
        assert "runPolicy" in doc["spec"]
