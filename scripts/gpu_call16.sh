#!/bin/bash
set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 300 python -m pytest tests/test_gemm_gpu.py -m gpu -q -x --timeout=200 -k "fused or probe or tc_linear" > gpurun_out/pytest16.log 2>&1; echo "pytest rc=$?" >> gpurun_out/rc16.txt
timeout 200 $TR --nproc-per-node 2 --master-port 29591 benchmarks/gemm_allreduce_bench.py --out gpurun_out/gemm_allreduce_2.json > gpurun_out/gemm_allreduce_2.log 2>&1; echo "gar2 rc=$?" >> gpurun_out/rc16.txt
cat gpurun_out/rc16.txt; tail -n 3 gpurun_out/pytest16.log; grep -E "^comm|^[0-9]+x|^  " gpurun_out/gemm_allreduce_2.log | head -60
