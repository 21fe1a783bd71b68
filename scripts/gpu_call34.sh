#!/bin/bash
# 2 GPUs: BERT-large TFJob (Chief + 1 Worker), NVLS vs P2P gradient all-reduce
set -x
mkdir -p gpurun_out
timeout 300 python examples/run_job.py examples/bert_large_dp/tfjob.yaml --replicas Worker=1 --arg=--steps=10 --out gpurun_out/job_bert_2_nvls.json > gpurun_out/job_bert_2_nvls.log 2>&1; echo "nvls rc=$?" > gpurun_out/rc34.txt
timeout 300 python examples/run_job.py examples/bert_large_dp/tfjob.yaml --replicas Worker=1 --arg=--steps=10 --arg=--no-nvls --out gpurun_out/job_bert_2_p2p.json > gpurun_out/job_bert_2_p2p.log 2>&1; echo "p2p rc=$?" >> gpurun_out/rc34.txt
cat gpurun_out/rc34.txt
python - <<'PY'
import json
for f in ['job_bert_2_nvls','job_bert_2_p2p']:
    try:
        d=json.load(open(f'gpurun_out/{f}.json')); print(f, d['result'])
        for pod,log in list(d['logs'].items())[:1]: print('   ', pod, [l[:500] for l in log.strip().splitlines() if l.startswith('{')][-1:])
    except Exception as e: print(f,'ERR',e)
PY
tail -n 5 gpurun_out/job_bert_2_nvls.log
