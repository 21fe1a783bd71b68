#!/bin/bash
# 2-GPU pass: real P2P all-reduce vs NCCL, 2-GPU bench (native + stock), kernel microbench
set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 300 python benchmarks/kernel_bench.py --out gpurun_out/kernel_bench.json > gpurun_out/kernel_bench.log 2>&1; echo "kernel_bench rc=$?" >> gpurun_out/rc2.txt
timeout 300 $TR --master-port 29511 benchmarks/allreduce_bench.py --out gpurun_out/allreduce_2.json > gpurun_out/allreduce_2.log 2>&1; echo "allreduce2 rc=$?" >> gpurun_out/rc2.txt
timeout 400 $TR --master-port 29512 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/bench2_native.json 2> gpurun_out/bench2_native.err; echo "bench2 native rc=$?" >> gpurun_out/rc2.txt
timeout 400 $TR --master-port 29513 bench.py --gpus 2 --steps 20 --warmup 5 --no-overlap > gpurun_out/bench2_noov.json 2> gpurun_out/bench2_noov.err; echo "bench2 no-overlap rc=$?" >> gpurun_out/rc2.txt
timeout 400 $TR --master-port 29514 bench.py --gpus 2 --steps 20 --warmup 5 --impl stock > gpurun_out/bench2_stock.json 2> gpurun_out/bench2_stock.err; echo "bench2 stock rc=$?" >> gpurun_out/rc2.txt
cat gpurun_out/rc2.txt
tail -n 3 gpurun_out/kernel_bench.log gpurun_out/allreduce_2.log
tail -n 5 gpurun_out/bench2_native.err gpurun_out/bench2_noov.err
cat gpurun_out/bench2_*.json
