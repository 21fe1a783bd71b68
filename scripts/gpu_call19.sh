#!/bin/bash
set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 300 python -m pytest tests/test_gemm_gpu.py -m gpu -q -x --timeout=200 -k "transpose or tc_linear or fused" > gpurun_out/pytest19.log 2>&1; echo "pytest rc=$?" >> gpurun_out/rc19.txt
timeout 200 $TR --nproc-per-node 2 --master-port 29611 benchmarks/gemm_allreduce_bench.py --out gpurun_out/gemm_allreduce_2b.json > gpurun_out/gemm_allreduce_2b.log 2>&1; echo "gar2 rc=$?" >> gpurun_out/rc19.txt
timeout 200 python examples/bert_large_dp/train.py --steps 5 --warmup 2 > gpurun_out/bert1_graph2.log 2>&1; echo "bert rc=$?" >> gpurun_out/rc19.txt
cat gpurun_out/rc19.txt; tail -n 3 gpurun_out/pytest19.log; grep -E "^comm|^[0-9]+x|^  " gpurun_out/gemm_allreduce_2b.log | head -50; tail -n 1 gpurun_out/bert1_graph2.log
