#!/bin/bash
# first GPU pass: numerics of every kernel, then 1-GPU bench variants
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit --format=csv > gpurun_out/smi.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -q -x --timeout=300 -k "not gemm" > gpurun_out/pytest_core.log 2>&1; echo "core rc=$?" >> gpurun_out/rc.txt
timeout 600 python -m pytest tests/test_gemm_gpu.py -m gpu -q --timeout=300 > gpurun_out/pytest_gemm.log 2>&1; echo "gemm rc=$?" >> gpurun_out/rc.txt
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_native.json 2> gpurun_out/bench_native.err; echo "bench native rc=$?" >> gpurun_out/rc.txt
timeout 600 python bench.py --steps 10 --warmup 3 --impl stock > gpurun_out/bench_stock.json 2> gpurun_out/bench_stock.err; echo "bench stock rc=$?" >> gpurun_out/rc.txt
timeout 600 python bench.py --steps 10 --warmup 3 --no-fused-bn > gpurun_out/bench_nofbn.json 2> gpurun_out/bench_nofbn.err; echo "bench nofbn rc=$?" >> gpurun_out/rc.txt
timeout 600 python bench.py --steps 10 --warmup 3 --no-graph > gpurun_out/bench_nograph.json 2> gpurun_out/bench_nograph.err; echo "bench nograph rc=$?" >> gpurun_out/rc.txt
cat gpurun_out/rc.txt
tail -5 gpurun_out/pytest_core.log gpurun_out/pytest_gemm.log
cat gpurun_out/bench_*.json
