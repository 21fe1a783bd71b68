#!/bin/bash
set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 150 $TR --nproc-per-node 2 --master-port 29571 benchmarks/gar_debug.py > gpurun_out/gar_debug.log 2>&1; echo "gar debug rc=$?" >> gpurun_out/rc14.txt
cat gpurun_out/rc14.txt; grep "^\[" gpurun_out/gar_debug.log | head -40
