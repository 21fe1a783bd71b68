#!/bin/bash
set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 300 $TR --nproc-per-node 8 --master-port 29541 benchmarks/allreduce_bench.py --out gpurun_out/allreduce_8_nvls.json > gpurun_out/allreduce_8_nvls.log 2>&1; echo "allreduce8 rc=$?" >> gpurun_out/rc10.txt
cat gpurun_out/rc10.txt; tail -n 4 gpurun_out/allreduce_8_nvls.log | cut -c1-600
