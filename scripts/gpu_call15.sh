#!/bin/bash
set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 150 $TR --nproc-per-node 2 --master-port 29581 benchmarks/gemm_allreduce_bench.py --out gpurun_out/gemm_allreduce_2.json > gpurun_out/gemm_allreduce_2.log 2>&1; echo "gar2 rc=$?" >> gpurun_out/rc15.txt
cat gpurun_out/rc15.txt; grep -E "^comm|^[0-9]+x|^  " gpurun_out/gemm_allreduce_2.log | head -60
