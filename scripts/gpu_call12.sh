#!/bin/bash
# 8-GPU: bench with NVLS, then the four example jobs at their BASELINE.json shapes through the launcher
set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 400 $TR --nproc-per-node 8 --master-port 29551 bench.py --gpus 8 --steps 30 --warmup 5 > gpurun_out/bench8_nvls.json 2> gpurun_out/bench8_nvls.err; echo "bench8 nvls rc=$?" >> gpurun_out/rc12.txt
timeout 400 $TR --nproc-per-node 8 --master-port 29552 bench.py --gpus 8 --steps 30 --warmup 5 --no-nvls > gpurun_out/bench8_p2p.json 2> gpurun_out/bench8_p2p.err; echo "bench8 p2p rc=$?" >> gpurun_out/rc12.txt
timeout 500 python examples/run_job.py examples/resnet50_ddp/pytorchjob.yaml --arg=--steps=30 --out gpurun_out/job_resnet_8.json > gpurun_out/job_resnet_8.log 2>&1; echo "resnet rc=$?" >> gpurun_out/rc12.txt
timeout 500 python examples/run_job.py examples/xgb_hist/xgboostjob.yaml --arg=--trees=3 --out gpurun_out/job_xgb_8.json > gpurun_out/job_xgb_8.log 2>&1; echo "xgb rc=$?" >> gpurun_out/rc12.txt
timeout 500 python examples/run_job.py examples/xgb_hist/xgboostjob.yaml --arg=--trees=3 --arg=--impl=nccl --out gpurun_out/job_xgb_8_nccl.json > gpurun_out/job_xgb_8_nccl.log 2>&1; echo "xgb nccl rc=$?" >> gpurun_out/rc12.txt
timeout 600 python examples/run_job.py examples/wide_deep_ps/tfjob.yaml --arg=--steps=30 --out gpurun_out/job_wd_8.json > gpurun_out/job_wd_8.log 2>&1; echo "widedeep rc=$?" >> gpurun_out/rc12.txt
timeout 700 python examples/run_job.py examples/bert_large_dp/tfjob.yaml --arg=--steps=10 --out gpurun_out/job_bert_8.json > gpurun_out/job_bert_8.log 2>&1; echo "bert rc=$?" >> gpurun_out/rc12.txt
timeout 700 python examples/run_job.py examples/bert_large_dp/tfjob.yaml --arg=--steps=10 --arg=--impl=nccl --out gpurun_out/job_bert_8_nccl.json > gpurun_out/job_bert_8_nccl.log 2>&1; echo "bert nccl rc=$?" >> gpurun_out/rc12.txt
cat gpurun_out/rc12.txt
cat gpurun_out/bench8_nvls.json gpurun_out/bench8_p2p.json
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/job_*_8*.json')):
    d=json.load(open(f)); print(f, d['result'])
    for pod,log in list(d['logs'].items())[:3]: print('   ', pod, [l[:300] for l in log.strip().splitlines()[-2:]])
PY
