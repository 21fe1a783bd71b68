#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 700 python -m pytest tests/test_gemm_gpu.py tests/test_layernorm_gpu.py -m gpu -q -x --timeout=300 > gpurun_out/pytest24.log 2>&1; echo "pytest rc=$?" > gpurun_out/rc24.txt
if grep -q "pytest rc=0" gpurun_out/rc24.txt; then
timeout 200 python benchmarks/kernel_bench.py --only gemm --out gpurun_out/gemm_v4.json > gpurun_out/gemm_v4.log 2>&1; echo "gemm rc=$?" >> gpurun_out/rc24.txt
timeout 300 python benchmarks/bert_breakdown.py --impls native --top 12 --out gpurun_out/bert_breakdown5.json > gpurun_out/bert_breakdown5.log 2>&1; echo "breakdown rc=$?" >> gpurun_out/rc24.txt
timeout 200 python examples/bert_large_dp/train.py --steps 5 --warmup 3 > gpurun_out/bert1_v6.log 2>&1; echo "bert rc=$?" >> gpurun_out/rc24.txt
fi
cat gpurun_out/rc24.txt; tail -n 25 gpurun_out/pytest24.log; grep "^{" gpurun_out/gemm_v4.log; grep -v Warn gpurun_out/bert_breakdown5.log | cut -c1-150 | head -34; tail -n 1 gpurun_out/bert1_v6.log
