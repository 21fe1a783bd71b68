#!/bin/bash
set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 600 python -m pytest tests -m gpu -q -x --timeout=300 -k "ps_ or noalloc" > gpurun_out/pytest9.log 2>&1; echo "pytest rc=$?" >> gpurun_out/rc9.txt
timeout 300 $TR --nproc-per-node 2 --master-port 29531 benchmarks/allreduce_bench.py --max-mb 64 --out gpurun_out/allreduce_2_nvls.json > gpurun_out/allreduce_2_nvls.log 2>&1; echo "allreduce2 rc=$?" >> gpurun_out/rc9.txt
cat gpurun_out/rc9.txt; tail -n 5 gpurun_out/pytest9.log; tail -n 12 gpurun_out/allreduce_2_nvls.log
