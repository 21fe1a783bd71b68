#!/bin/bash
# 8 GPUs: BERT-large TFJob (Chief + 7 Workers) through the operator + launcher, native path
set -x
mkdir -p gpurun_out
timeout 400 python examples/run_job.py examples/bert_large_dp/tfjob.yaml --arg=--steps=12 --out gpurun_out/job_bert_8_v2.json > gpurun_out/job_bert_8_v2.log 2>&1; echo "bert8 rc=$?" > gpurun_out/rc29.txt
cat gpurun_out/rc29.txt
python - <<'PY'
import json
d=json.load(open('gpurun_out/job_bert_8_v2.json')); print(d['result'])
for pod,log in d['logs'].items():
    print(pod, [l[:400] for l in log.strip().splitlines() if l.startswith('{')][-1:])
PY
