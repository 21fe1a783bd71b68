"""Deployment manifests + generated CRDs (manifests/, api/crd.py)."""
import glob
import os

import yaml

from tf_operator_b200.api import KINDS, crd

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_all_manifests_parse_and_kustomizations_resolve():
    for k in glob.glob(os.path.join(ROOT, "manifests", "**", "kustomization.yaml"), recursive=True):
        doc = yaml.safe_load(open(k))
        for res in doc["resources"]:
            assert os.path.exists(os.path.join(os.path.dirname(k), res)), (k, res)
    for f in glob.glob(os.path.join(ROOT, "manifests", "**", "*.yaml"), recursive=True):
        assert list(yaml.safe_load_all(open(f)))


def test_generated_crds_are_committed_and_match_the_contract():
    for mod in KINDS.values():
        i = mod.INFO
        want = crd.crd_for(mod)
        have = yaml.safe_load(open(os.path.join(ROOT, "manifests", "base", f"kubeflow.org_{i.plural}.yaml")))
        assert have == want, f"regenerate with python -m tf_operator_b200.api.crd ({i.kind})"
        assert have["metadata"]["name"] == f"{i.plural}.kubeflow.org"
        v = have["spec"]["versions"][0]
        assert v["name"] == "v1" and v["served"] and v["storage"] and v["subresources"] == {"status": {}}
        spec = v["schema"]["openAPIV3Schema"]["properties"]["spec"]
        assert i.specs_key in spec["required"]
        assert set(spec["properties"]["runPolicy"]["properties"]) == {
            "activeDeadlineSeconds", "backoffLimit", "cleanPodPolicy", "schedulingPolicy", "ttlSecondsAfterFinished"}
        st = v["schema"]["openAPIV3Schema"]["properties"]["status"]
        assert st["required"] == ["conditions", "replicaStatuses"]
    mx = yaml.safe_load(open(os.path.join(ROOT, "manifests", "base", "kubeflow.org_mxjobs.yaml")))
    assert "jobMode" in mx["spec"]["versions"][0]["schema"]["openAPIV3Schema"]["properties"]["spec"]["required"]


def test_example_job_manifests_validate():
    from tf_operator_b200 import api
    for f in glob.glob(os.path.join(ROOT, "examples", "**", "*.yaml"), recursive=True):
        job = yaml.safe_load(open(f))
        if job.get("kind") == "CustomResourceDefinition":     # examples/crd: legacy CRD manifests
            assert job["spec"]["group"] in ("kubeflow.org", "scheduling.volcano.sh",
                                            "scheduling.incubator.k8s.io"), f
            continue
        api.set_defaults(job)
        api.validate(job)
