#!/bin/bash
# 2-GPU: fused GEMM+all-reduce (emulated test + real), BERT graph path, xgb
set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 600 python -m pytest tests/test_gemm_gpu.py -m gpu -q -x --timeout=300 > gpurun_out/pytest13.log 2>&1; echo "pytest gemm rc=$?" >> gpurun_out/rc13.txt
timeout 300 $TR --nproc-per-node 2 --master-port 29561 benchmarks/gemm_allreduce_bench.py --out gpurun_out/gemm_allreduce_2.json > gpurun_out/gemm_allreduce_2.log 2>&1; echo "gar2 rc=$?" >> gpurun_out/rc13.txt
timeout 300 python benchmarks/kernel_bench.py --out gpurun_out/kernel_bench13.json > gpurun_out/kernel_bench13.log 2>&1; echo "kb rc=$?" >> gpurun_out/rc13.txt
timeout 400 python examples/run_job.py examples/bert_large_dp/tfjob.yaml --replicas Worker=1 --arg=--model=base --arg=--steps=5 --arg=--batch-size=8 --out gpurun_out/job_bert_2g.json > gpurun_out/job_bert_2g.log 2>&1; echo "bert graph rc=$?" >> gpurun_out/rc13.txt
timeout 400 python examples/run_job.py examples/xgb_hist/xgboostjob.yaml --replicas Worker=1 --arg=--rows=200000 --arg=--trees=3 --out gpurun_out/job_xgb_2.json > gpurun_out/job_xgb_2.log 2>&1; echo "xgb rc=$?" >> gpurun_out/rc13.txt
cat gpurun_out/rc13.txt; tail -n 5 gpurun_out/pytest13.log; tail -n 3 gpurun_out/gemm_allreduce_2.log | cut -c1-1500
python - <<'PY'
import json,glob
for f in ['gpurun_out/job_bert_2g.json','gpurun_out/job_xgb_2.json']:
    try:
        d=json.load(open(f)); print(f, d['result'])
        for pod,log in list(d['logs'].items())[:2]: print('   ', pod, [l[:500] for l in log.strip().splitlines()[-2:]])
    except Exception as e: print(f, 'ERR', e)
k=json.load(open('gpurun_out/kernel_bench13.json'))
for r in k['gemm']: print(r)
PY
