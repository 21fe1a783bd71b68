"""The SDK under the reference's import path: ``from kubeflow.tfjob import TFJobClient, V1TFJob``
(sdk/python/kubeflow/tfjob/__init__.py:17-36) and the reference's own SDK e2e test
(sdk/python/test/test_e2e.py:30-72) executed UNMODIFIED against the single-box cluster."""
import importlib
import os
import sys
import types

import pytest

REF_TEST = "/root/reference/sdk/python/test/test_e2e.py"
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_import_paths_resolve_to_the_sdk():
    import kubeflow.tfjob as kt
    from tf_operator_b200 import sdk
    assert kt.TFJobClient is sdk.TFJobClient and kt.V1TFJob is sdk.V1TFJob
    for mod, name in [("kubeflow.tfjob.api.tf_job_client", "TFJobClient"),
                      ("kubeflow.tfjob.api.tf_job_watch", "watch"),
                      ("kubeflow.tfjob.constants.constants", "TFJOB_PLURAL"),
                      ("kubeflow.tfjob.utils.utils", "get_labels"),
                      ("kubeflow.tfjob.utils.utils", "set_tfjob_namespace"),
                      ("kubeflow.tfjob.models.v1_tf_job_spec", "V1TFJobSpec"),
                      ("kubeflow.tfjob.models.v1_job_condition", "V1JobCondition"),
                      ("kubeflow.tfjob.models.v1_time", "V1Time"),
                      ("kubeflow.tfjob.api_client", "ApiClient"),
                      ("kubeflow.tfjob.configuration", "Configuration"),
                      ("kubeflow.tfjob.rest", "ApiException"),
                      ("kubeflow.pytorchjob", "PyTorchJobClient")]:
        assert hasattr(importlib.import_module(mod), name), (mod, name)
    # the flat pre-RunPolicy spec of the reference's swagger model
    spec = kt.V1TFJobSpec(clean_pod_policy="None", tf_replica_specs={})
    assert spec.to_dict()["clean_pod_policy"] == "None"


def test_pyproject_declares_both_packages():
    import tomllib
    with open(os.path.join(REPO, "pyproject.toml"), "rb") as f:
        meta = tomllib.load(f)
    inc = meta["tool"]["setuptools"]["packages"]["find"]["include"]
    assert "kubeflow*" in inc and "tf_operator_b200*" in inc
    assert meta["project"]["scripts"]["tfopctl"] == "tf_operator_b200.cli:main"


@pytest.mark.skipif(not os.path.exists(REF_TEST), reason="reference tree not present")
def test_reference_sdk_e2e_runs_unmodified(tmp_path, monkeypatch):
    """Executes the reference file's source as is. The image it names becomes a directory of the
    local image registry (KubeletOptions.image_roots) holding our port of the same script."""
    from tf_operator_b200 import cluster as cl
    from tf_operator_b200.launcher.kubelet import KubeletOptions
    from tf_operator_b200.sdk import compat

    try:
        import kubernetes  # noqa: F401
    except ImportError:
        monkeypatch.syspath_prepend(compat.PATH)
    root = tmp_path / "image"
    (root / "var" / "tf_mnist").mkdir(parents=True)
    (root / "train").mkdir()
    os.symlink(os.path.join(REPO, "examples", "mnist_with_summaries", "mnist_with_summaries.py"),
               root / "var" / "tf_mnist" / "mnist_with_summaries.py")
    opts = KubeletOptions(log_dir=str(tmp_path / "logs"),
                          image_roots={"gcr.io/kubeflow-ci/tf-mnist-with-summaries:1.0": str(root)},
                          )
    c = cl.LocalCluster(kubelet_options=opts).start()
    cl.reset_default_cluster()
    monkeypatch.setattr(cl, "_default", c, raising=False)
    try:
        src = open(REF_TEST).read()
        mod = types.ModuleType("ref_sdk_e2e")
        mod.__file__ = REF_TEST
        exec(compile(src, REF_TEST, "exec"), mod.__dict__)
        # shorten the reference's 30 s poll (tf_job_client.py:223-228 defaults) - same calls
        orig = mod.TFJOB_CLIENT.wait_for_job
        monkeypatch.setattr(mod.TFJOB_CLIENT, "wait_for_job",
                            lambda name, **kw: orig(name, polling_interval=0.5,
                                                    timeout_seconds=240, **kw))
        mod.test_sdk_e2e()
        with pytest.raises(RuntimeError):
            mod.TFJOB_CLIENT.get("mnist-ci-test", namespace="kubeflow")
        assert (root / "train" / "logs").exists()
    finally:
        c.stop()
        for k in [k for k in sys.modules if k == "kubernetes" or k.startswith("kubernetes.")]:
            if compat.PATH in (getattr(sys.modules[k], "__file__", "") or ""):
                del sys.modules[k]
