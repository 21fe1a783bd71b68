#!/bin/bash
# 8-GPU: fused GEMM+all-reduce bench, BERT-large native (graph) vs all-library baseline, xgb hist timing
set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 200 $TR --nproc-per-node 8 --master-port 29601 benchmarks/gemm_allreduce_bench.py --out gpurun_out/gemm_allreduce_8.json > gpurun_out/gemm_allreduce_8.log 2>&1; echo "gar8 rc=$?" >> gpurun_out/rc17.txt
timeout 400 python examples/run_job.py examples/bert_large_dp/tfjob.yaml --arg=--steps=15 --out gpurun_out/job_bert_8_graph.json > gpurun_out/job_bert_8_graph.log 2>&1; echo "bert graph rc=$?" >> gpurun_out/rc17.txt
timeout 400 python examples/run_job.py examples/bert_large_dp/tfjob.yaml --arg=--steps=15 --arg=--impl=nccl --out gpurun_out/job_bert_8_stock.json > gpurun_out/job_bert_8_stock.log 2>&1; echo "bert stock rc=$?" >> gpurun_out/rc17.txt
timeout 300 python examples/run_job.py examples/xgb_hist/xgboostjob.yaml --arg=--trees=4 --out gpurun_out/job_xgb_8b.json > gpurun_out/job_xgb_8b.log 2>&1; echo "xgb rc=$?" >> gpurun_out/rc17.txt
timeout 300 python examples/run_job.py examples/xgb_hist/xgboostjob.yaml --arg=--trees=4 --arg=--impl=nccl --out gpurun_out/job_xgb_8b_nccl.json > gpurun_out/job_xgb_8b_nccl.log 2>&1; echo "xgb nccl rc=$?" >> gpurun_out/rc17.txt
cat gpurun_out/rc17.txt; grep -E "^comm|^[0-9]+x|^  " gpurun_out/gemm_allreduce_8.log | head -40
python - <<'PY'
import json
for f in ['job_bert_8_graph','job_bert_8_stock','job_xgb_8b','job_xgb_8b_nccl']:
    try:
        d=json.load(open(f'gpurun_out/{f}.json')); print(f, d['result'])
        for pod,log in list(d['logs'].items())[:1]: print('   ', pod, [l[:600] for l in log.strip().splitlines() if l.startswith('{')][-1:])
    except Exception as e: print(f,'ERR',e)
PY
