import json, os, sys, time, yaml
REPO = os.environ.get("GRAFT_REPO_ROOT", "/root/repo")
sys.path.insert(0, REPO)
from tf_operator_b200.cluster import LocalCluster
from tf_operator_b200.launcher.kubelet import KubeletOptions
from tf_operator_b200.sdk import TFJobClient
out = {}
cl = LocalCluster(kubelet_options=KubeletOptions(cwd=REPO))
try:
    cli = TFJobClient(cluster=cl)
    for mode, extra in (("native_bucket", []), ("native_fused_wgrad", ["--fused-wgrad"])):
        job = yaml.safe_load(open(os.path.join(REPO, "examples", "bert_large_dp", "tfjob.yaml")))
        name = job["metadata"]["name"] = "bert-" + mode.replace("_", "-")
        specs = job["spec"]["tfReplicaSpecs"]
        for rt in list(specs):
            ctr = specs[rt]["template"]["spec"]["containers"][0]
            ctr["command"] = [c for c in ctr["command"] if not str(c).startswith("--ckpt")] + ["--steps=20", "--warmup=4"] + extra
        t0 = time.time()
        cli.create(job)
        cli.wait_for_job(name, polling_interval=0.2, timeout_seconds=600)
        logs = cli.get_logs(name, master=False)
        rows = [json.loads(l) for t in logs.values() for l in t.splitlines() if l.startswith('{"role"')]
        if not cli.is_job_succeeded(name) or not rows:
            out[mode] = {"error": "\n".join(f"--- {k}\n{v[-1200:]}" for k, v in logs.items())}
        else:
            out[mode] = {"ms_per_step": max(r["ms_per_step"] for r in rows), "seq_per_s_job": min(r["seq_per_s_job"] for r in rows),
                         "fused_wgrad": rows[0].get("fused_wgrad"), "allreduce": rows[0]["allreduce"], "loss": rows[0]["loss"],
                         "kernel_launches_per_step": rows[0]["kernel_launches_per_step"], "world": rows[0]["world"],
                         "job_wall_s": round(time.time() - t0, 1)}
        cli.delete(name)
        time.sleep(2)
finally:
    cl.stop()
print(json.dumps(out, indent=1))
json.dump(out, open(os.path.join(REPO, "gpurun_out", "bert_8gpu.json"), "w"), indent=1)
