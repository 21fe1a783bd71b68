#!/bin/bash
# 8-GPU box: scaling of bench.py at N=4,8 (native + stock at 8), all-reduce microbench at 4 and 8
set -x
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo.txt 2>&1
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 300 $TR --nproc-per-node 8 --master-port 29521 benchmarks/allreduce_bench.py --out gpurun_out/allreduce_8.json > gpurun_out/allreduce_8.log 2>&1; echo "allreduce8 rc=$?" >> gpurun_out/rc8.txt
timeout 300 $TR --nproc-per-node 4 --master-port 29522 benchmarks/allreduce_bench.py --out gpurun_out/allreduce_4.json > gpurun_out/allreduce_4.log 2>&1; echo "allreduce4 rc=$?" >> gpurun_out/rc8.txt
timeout 400 $TR --nproc-per-node 8 --master-port 29523 bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/bench8_native.json 2> gpurun_out/bench8_native.err; echo "bench8 native rc=$?" >> gpurun_out/rc8.txt
timeout 400 $TR --nproc-per-node 4 --master-port 29524 bench.py --gpus 4 --steps 20 --warmup 5 > gpurun_out/bench4_native.json 2> gpurun_out/bench4_native.err; echo "bench4 native rc=$?" >> gpurun_out/rc8.txt
timeout 400 $TR --nproc-per-node 8 --master-port 29525 bench.py --gpus 8 --steps 20 --warmup 5 --impl stock > gpurun_out/bench8_stock.json 2> gpurun_out/bench8_stock.err; echo "bench8 stock rc=$?" >> gpurun_out/rc8.txt
NCCL_DEBUG=INFO timeout 200 $TR --nproc-per-node 8 --master-port 29526 benchmarks/allreduce_bench.py --max-mb 16 2>&1 | grep -i -E "nvls|NVLS|multicast" | head -20 > gpurun_out/nccl_nvls.txt
cat gpurun_out/rc8.txt; tail -n 5 gpurun_out/bench8_native.err; cat gpurun_out/bench8_native.json gpurun_out/bench4_native.json gpurun_out/bench8_stock.json
