#!/bin/bash
# 8 GPUs: BERT-large TFJob (NVLS gradient all-reduce) and the ResNet-50 bench at N=8
set -x
mkdir -p gpurun_out
timeout 300 python examples/run_job.py examples/bert_large_dp/tfjob.yaml --arg=--steps=12 --out gpurun_out/job_bert_8_v3.json > gpurun_out/job_bert_8_v3.log 2>&1; echo "bert8 rc=$?" > gpurun_out/rc36.txt
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29655 bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/bench8_v2.json 2> gpurun_out/bench8_v2.err; echo "bench8 rc=$?" >> gpurun_out/rc36.txt
cat gpurun_out/rc36.txt
python - <<'PY'
import json
d=json.load(open('gpurun_out/job_bert_8_v3.json')); print(d['result'])
for pod,log in list(d['logs'].items())[:2]:
    print(pod, [l[:460] for l in log.strip().splitlines() if l.startswith('{')][-1:])
PY
cut -c1-700 gpurun_out/bench8_v2.json; tail -n 3 gpurun_out/bench8_v2.err
