#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_layernorm_gpu.py tests/test_gemm_gpu.py -m gpu -q -x --timeout=200 > gpurun_out/pytest23.log 2>&1; echo "pytest rc=$?" > gpurun_out/rc23.txt
timeout 300 python benchmarks/bert_breakdown.py --impls native --top 22 --out gpurun_out/bert_breakdown4.json > gpurun_out/bert_breakdown4.log 2>&1; echo "breakdown rc=$?" >> gpurun_out/rc23.txt
timeout 200 python examples/bert_large_dp/train.py --steps 5 --warmup 3 > gpurun_out/bert1_v5.log 2>&1; echo "bert rc=$?" >> gpurun_out/rc23.txt
cat gpurun_out/rc23.txt; tail -n 15 gpurun_out/pytest23.log; grep -v Warn gpurun_out/bert_breakdown4.log | cut -c1-150 | head -50; tail -n 1 gpurun_out/bert1_v5.log
