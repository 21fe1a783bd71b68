"""tfopctl (tf_operator_b200/cli.py): the kubectl verbs of the reference's quick-start, against
the REST facade of a running single-box cluster."""
import os
import sys

import pytest
import yaml

from tf_operator_b200 import cli
from tf_operator_b200.cluster import LocalCluster
from tf_operator_b200.launcher.kubelet import KubeletOptions
from tf_operator_b200.launcher.topology import synthetic

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TEST_APP = [sys.executable, os.path.join(REPO, "examples", "test_server", "test_app.py")]


@pytest.fixture()
def cluster():
    cl = LocalCluster(topology=synthetic(0), serve_http=True,
                      kubelet_options=KubeletOptions(cwd=REPO, restart_backoff=0.05))
    yield cl
    cl.stop()


def run(cluster, capsys, *argv):
    rc = cli.main(["--server", cluster.api.url, *argv])
    out = capsys.readouterr()
    return rc, out.out, out.err


def test_apply_get_describe_logs_wait_delete(cluster, capsys, tmp_path):
    job = {"apiVersion": "kubeflow.org/v1", "kind": "TFJob", "metadata": {"name": "ctl"},
           "spec": {"tfReplicaSpecs": {"Worker": {"replicas": 2, "restartPolicy": "Never", "template": {
               "spec": {"containers": [{"name": "tensorflow", "image": "x",
                                        "command": ["sh", "-c", "echo hello from $TFOP_POD_NAME"]}]}}}}}}
    f = tmp_path / "job.yaml"
    f.write_text(yaml.safe_dump(job))
    rc, out, _ = run(cluster, capsys, "apply", "-f", str(f))
    assert rc == 0 and "tfjob.kubeflow.org/ctl created" in out
    rc, out, _ = run(cluster, capsys, "wait", "tfjob", "ctl", "--for", "Succeeded", "--timeout", "30")
    assert rc == 0 and "condition met: Succeeded" in out
    rc, out, _ = run(cluster, capsys, "get", "tfjobs")
    assert rc == 0 and out.splitlines()[0].split() == ["NAME", "STATE", "AGE"]
    assert out.splitlines()[1].split()[:2] == ["ctl", "Succeeded"]
    rc, out, _ = run(cluster, capsys, "get", "pods", "-l", "job-name=ctl", "-o", "name")
    assert sorted(out.split()) == ["pod/ctl-worker-0", "pod/ctl-worker-1"]
    rc, out, _ = run(cluster, capsys, "logs", "ctl-worker-1")
    assert "hello from ctl-worker-1" in out
    rc, out, _ = run(cluster, capsys, "describe", "tfjob", "ctl")
    assert "Conditions:" in out and "TFJobSucceeded" in out and "Replica Statuses:" in out
    assert "Created pod: ctl-worker-0" in out
    rc, out, _ = run(cluster, capsys, "get", "tfjob", "ctl", "-o", "json")
    assert '"kind": "TFJob"' in out
    rc, out, _ = run(cluster, capsys, "apply", "-f", str(f))
    assert "configured" in out
    rc, out, _ = run(cluster, capsys, "get", "events")
    assert rc == 0 and "REASON" in out
    rc, out, _ = run(cluster, capsys, "delete", "tfjob", "ctl")
    assert rc == 0 and 'tfjob "ctl" deleted' in out
    rc, out, _ = run(cluster, capsys, "get", "tfjobs")
    assert "No resources found" in out


def test_errors_are_reported_like_kubectl(cluster, capsys, tmp_path):
    rc, _, err = run(cluster, capsys, "get", "tfjob", "nope")
    assert rc == 1 and "Error from server (NotFound)" in err
    bad = tmp_path / "bad.yaml"
    bad.write_text(yaml.safe_dump({"apiVersion": "kubeflow.org/v1", "kind": "TFJob",
                                   "metadata": {"name": "bad"}, "spec": {"notTheActualField": {}}}))
    rc, _, err = run(cluster, capsys, "apply", "-f", str(bad))
    assert rc == 1 and "tfReplicaSpecs" in err
    rc, _, err = run(cluster, capsys, "get", "widgets")
    assert rc == 1 and "unknown job kind" in err
    rc, _, err = run(cluster, capsys, "wait", "tfjob", "nope", "--timeout", "1")
    assert rc == 1
