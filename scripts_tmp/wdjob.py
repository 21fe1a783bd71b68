import json, os, sys, time, yaml
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
from tf_operator_b200.cluster import LocalCluster
from tf_operator_b200.launcher.kubelet import KubeletOptions
from tf_operator_b200.sdk import TFJobClient
job = yaml.safe_load(open(os.path.join(REPO, "examples", "wide_deep_ps", "tfjob.yaml")))
specs = job["spec"]["tfReplicaSpecs"]
specs["PS"]["replicas"], specs["Worker"]["replicas"] = 1, 1
for rt in specs:
    ctr = specs[rt]["template"]["spec"]["containers"][0]
    ctr["command"] = list(ctr["command"]) + ["--steps=30", "--profile"]
cl = LocalCluster(kubelet_options=KubeletOptions(cwd=REPO))
try:
    cli = TFJobClient(cluster=cl)
    cli.create(job)
    cli.wait_for_job("wide-deep-ps", polling_interval=0.2, timeout_seconds=600)
    for k, v in cli.get_logs("wide-deep-ps", master=False).items():
        print("=====", k)
        print("\n".join(l[:190] for l in v.splitlines() if not l.startswith("---")))
finally:
    cl.stop()
