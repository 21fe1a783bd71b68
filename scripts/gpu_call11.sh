#!/bin/bash
# 2-GPU: the example jobs through the launcher (SDK -> operator -> kubelet -> replicas)
set -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests -m gpu -q -x --timeout=300 -k "hist or ps_" > gpurun_out/pytest11.log 2>&1; echo "pytest rc=$?" >> gpurun_out/rc11.txt
timeout 400 python examples/run_job.py examples/resnet50_ddp/pytorchjob.yaml --replicas Worker=1 --arg=--steps=20 --out gpurun_out/job_resnet_2.json > gpurun_out/job_resnet_2.log 2>&1; echo "resnet rc=$?" >> gpurun_out/rc11.txt
timeout 400 python examples/run_job.py examples/bert_large_dp/tfjob.yaml --replicas Worker=1 --arg=--model=base --arg=--steps=5 --arg=--batch-size=8 --out gpurun_out/job_bert_2.json > gpurun_out/job_bert_2.log 2>&1; echo "bert rc=$?" >> gpurun_out/rc11.txt
timeout 400 python examples/run_job.py examples/wide_deep_ps/tfjob.yaml --replicas PS=1 --replicas Worker=1 --arg=--rows=1000000 --arg=--steps=10 --out gpurun_out/job_wd_2.json > gpurun_out/job_wd_2.log 2>&1; echo "widedeep rc=$?" >> gpurun_out/rc11.txt
timeout 400 python examples/run_job.py examples/xgb_hist/xgboostjob.yaml --replicas Worker=1 --arg=--rows=200000 --arg=--trees=2 --out gpurun_out/job_xgb_2.json > gpurun_out/job_xgb_2.log 2>&1; echo "xgb rc=$?" >> gpurun_out/rc11.txt
cat gpurun_out/rc11.txt; tail -n 3 gpurun_out/pytest11.log; tail -n 4 gpurun_out/job_*_2.log
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/job_*_2.json')):
    d=json.load(open(f)); print(f, d['result'])
    for pod,log in d['logs'].items(): print('   ', pod, log.strip().splitlines()[-3:])
PY
