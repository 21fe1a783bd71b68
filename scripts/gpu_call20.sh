#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_gemm_gpu.py -m gpu -q -x --timeout=200 > gpurun_out/pytest20.log 2>&1; echo "pytest rc=$?" > gpurun_out/rc20.txt
timeout 200 python benchmarks/kernel_bench.py --only gemm --out gpurun_out/gemm_v3.json > gpurun_out/gemm_v3.log 2>&1; echo "gemm rc=$?" >> gpurun_out/rc20.txt
timeout 300 python benchmarks/bert_breakdown.py --out gpurun_out/bert_breakdown.json > gpurun_out/bert_breakdown.log 2>&1; echo "breakdown rc=$?" >> gpurun_out/rc20.txt
timeout 200 python examples/bert_large_dp/train.py --steps 5 --warmup 3 > gpurun_out/bert1_v3.log 2>&1; echo "bert rc=$?" >> gpurun_out/rc20.txt
cat gpurun_out/rc20.txt; tail -n 3 gpurun_out/pytest20.log; grep "^{" gpurun_out/gemm_v3.log; tail -n 1 gpurun_out/bert1_v3.log
