#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 700 python -m pytest tests/test_layernorm_gpu.py -m gpu -q -x --timeout=300 > gpurun_out/pytest28.log 2>&1; echo "pytest rc=$?" > gpurun_out/rc28.txt
timeout 300 python benchmarks/bert_breakdown.py --impls native --top 16 --out gpurun_out/bert_breakdown7.json > gpurun_out/bert_breakdown7.log 2>&1; echo "breakdown rc=$?" >> gpurun_out/rc28.txt
timeout 200 python examples/bert_large_dp/train.py --steps 8 --warmup 3 > gpurun_out/bert1_v8.log 2>&1; echo "bert rc=$?" >> gpurun_out/rc28.txt
cat gpurun_out/rc28.txt; tail -n 15 gpurun_out/pytest28.log; grep -v Warn gpurun_out/bert_breakdown7.log | cut -c1-150 | grep -v "  gemm " | head -24; tail -n 1 gpurun_out/bert1_v8.log
