#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_gemm_gpu.py tests/test_kernels_gpu.py -m gpu -q -x --timeout=200 -k "tc_linear or adam or optim" > gpurun_out/pytest21.log 2>&1; echo "pytest rc=$?" > gpurun_out/rc21.txt
timeout 300 python benchmarks/bert_breakdown.py --impls native --top 12 --out gpurun_out/bert_breakdown2.json > gpurun_out/bert_breakdown2.log 2>&1; echo "breakdown rc=$?" >> gpurun_out/rc21.txt
timeout 300 python benchmarks/gemm_sustained.py --out gpurun_out/gemm_sustained.json > gpurun_out/gemm_sustained.log 2>&1; echo "sustained rc=$?" >> gpurun_out/rc21.txt
cat gpurun_out/rc21.txt; tail -n 3 gpurun_out/pytest21.log; grep -v Warn gpurun_out/bert_breakdown2.log | cut -c1-150 | head -40; grep "^{" gpurun_out/gemm_sustained.log
