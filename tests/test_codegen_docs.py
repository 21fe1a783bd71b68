"""Generated artefacts stay in sync with the API types (hack/verify-codegen.sh), the SDK docs
cover every exported client / model, and the setup script brings the operator up
(E10)."""
import os
import re
import signal
import socket
import subprocess
import sys

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_verify_codegen_passes():
    res = subprocess.run(["bash", "hack/verify-codegen.sh"], cwd=REPO, capture_output=True, text=True,
                         timeout=300)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-2000:]
    assert "up to date" in res.stdout


def test_sdk_docs_cover_every_client_and_model():
    from tf_operator_b200 import sdk
    from tf_operator_b200.sdk import models
    for client in ("TFJobClient", "PyTorchJobClient", "MXJobClient", "XGBoostJobClient"):
        text = open(os.path.join(REPO, "docs", "sdk", f"{client}.md")).read()
        for method in ("create", "get", "patch", "delete", "wait_for_job", "wait_for_condition",
                       "get_job_status", "is_job_running", "is_job_succeeded", "get_pod_names", "get_logs"):
            assert f"## {method}" in text, (client, method)
            assert hasattr(getattr(sdk, client), method)
        # every method row of the table carries a one-line description
        rows = [l for l in text.splitlines() if l.startswith("| [")]
        assert rows and all(not l.rstrip().endswith("|  |") for l in rows), client
    exported = [n for n in dir(sdk) if n.startswith("V1")]
    assert len(exported) >= 20
    for name in exported:
        page = open(os.path.join(REPO, "docs", "sdk", "models", f"{name}.md")).read()
        for attr, key in getattr(models, name).attribute_map.items():
            assert f"`{attr}`" in page and f"`{key}`" in page, (name, attr)
    api = open(os.path.join(REPO, "docs", "api", "kubeflow.org_v1_generated.md")).read()
    for kind, field in (("TFJob", "spec.tfReplicaSpecs"), ("PyTorchJob", "spec.pytorchReplicaSpecs"),
                        ("MXJob", "spec.jobMode"), ("XGBoostJob", "spec.xgbReplicaSpecs")):
        assert f"## {kind}" in api and f"`{field}`" in api


def test_setup_script_starts_the_operator_and_the_sdk_reaches_it(tmp_path):
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    env = dict(os.environ, TFOP_LOG=str(tmp_path / "operator.log"))
    res = subprocess.run(["bash", "scripts/setup-training-operator.sh", str(port)], cwd=REPO, env=env,
                         capture_output=True, text=True, timeout=120)
    pid = int(re.search(r"pid (\d+)", res.stdout).group(1))
    try:
        assert res.returncode == 0 and "ready:" in res.stdout, res.stdout + res.stderr
        probe = ("from tf_operator_b200.sdk import PyTorchJobClient; "
                 "print(PyTorchJobClient().get(namespace='default')['kind'])")
        out = subprocess.run([sys.executable, "-c", probe], cwd=REPO, capture_output=True, text=True,
                             env=dict(os.environ, TFOP_APISERVER=f"http://127.0.0.1:{port}"), timeout=60)
        assert out.stdout.strip() == "PyTorchJobList", out.stdout + out.stderr
    finally:
        os.kill(pid, signal.SIGTERM)


def test_py_checks_and_release_tooling(tmp_path):
    """hack/py_checks.py (every .py compiles, junit report) and hack/release.py (tarball +
    checksum manifest) - the reference's py_checks.py / release.py re-expressed for one box."""
    import json
    import tarfile
    from xml.etree import ElementTree as ET
    junit = tmp_path / "junit.xml"
    res = subprocess.run([sys.executable, "hack/py_checks.py", "--no-tests", "--junit_path", str(junit)],
                         cwd=REPO, capture_output=True, text=True, timeout=300)
    assert res.returncode == 0, res.stdout[-2000:]
    root = ET.parse(junit).getroot()
    assert int(root.attrib["tests"]) > 80 and root.attrib["failures"] == "0"
    out = tmp_path / "dist"
    res = subprocess.run([sys.executable, "hack/release.py", "--out", str(out), "--no-build"], cwd=REPO,
                         capture_output=True, text=True, timeout=300)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-2000:]
    meta = json.loads(res.stdout.strip().splitlines()[-1])
    tar = out / meta["artifact"]
    assert tar.exists() and meta["arch"] == "sm_100a" and len(meta["sha256"]) == 64
    with tarfile.open(tar) as t:
        names = t.getnames()
    assert any(n.endswith("tf_operator_b200/csrc/gemm_tcgen05.cu") for n in names)
    assert any(n.endswith("manifests/base/kubeflow.org_tfjobs.yaml") for n in names)
    assert not any("__pycache__" in n for n in names)
