#!/usr/bin/env python
"""The SDK walk-through of the reference notebook (sdk/python/examples/kubeflow-tfjob-sdk.ipynb):
define a TFJob with the model classes -> create -> get -> get_job_status -> wait_for_job(watch=True)
-> is_job_succeeded -> get_logs -> delete, on the single-box cluster.

    python examples/sdk/tfjob_sdk_walkthrough.py
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))

from tf_operator_b200.cluster import LocalCluster  # noqa: E402
from tf_operator_b200.launcher.kubelet import KubeletOptions  # noqa: E402
from tf_operator_b200.launcher.topology import synthetic  # noqa: E402
from tf_operator_b200.sdk import (TFJobClient, V1Container, V1ObjectMeta, V1PodSpec,  # noqa: E402
                                  V1PodTemplateSpec, V1ReplicaSpec, V1TFJob, V1TFJobSpec)

REPO = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main() -> int:
    namespace = "default"
    container = V1Container(name="tensorflow", image="tf-operator-b200/tf-smoke:local",
                            command=["python", "examples/tf_sample/tf_smoke.py"])
    worker = V1ReplicaSpec(replicas=2, restart_policy="Never",
                           template=V1PodTemplateSpec(spec=V1PodSpec(containers=[container])))
    chief = V1ReplicaSpec(replicas=1, restart_policy="Never",
                          template=V1PodTemplateSpec(spec=V1PodSpec(containers=[container])))
    ps = V1ReplicaSpec(replicas=1, restart_policy="Never",
                       template=V1PodTemplateSpec(spec=V1PodSpec(containers=[container])))
    tfjob = V1TFJob(api_version="kubeflow.org/v1", kind="TFJob",
                    metadata=V1ObjectMeta(name="mnist", namespace=namespace),
                    spec=V1TFJobSpec(clean_pod_policy="None",
                                     tf_replica_specs={"Worker": worker, "Chief": chief, "PS": ps}))
    with LocalCluster(topology=synthetic(0), kubelet_options=KubeletOptions(cwd=REPO)) as cluster:
        client = TFJobClient(cluster=cluster)
        created = client.create(tfjob, namespace=namespace)
        print("created", created["metadata"]["name"], "uid", created["metadata"]["uid"])
        print("get ->", client.get("mnist", namespace=namespace)["metadata"]["name"])
        print("status now:", client.get_job_status("mnist", namespace=namespace) or "(none yet)")
        client.wait_for_job("mnist", namespace=namespace, watch=True)
        ok = client.is_job_succeeded("mnist", namespace=namespace)
        print("succeeded:", ok)
        logs = client.get_logs("mnist", namespace=namespace)
        for pod, text in (logs or {}).items():
            print(f"--- {pod}\n" + "\n".join(text.strip().splitlines()[-3:]))
        client.delete("mnist", namespace=namespace)
        print("deleted")
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
