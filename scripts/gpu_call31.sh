#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gemm_gpu.py tests/test_engine_gpu.py -m gpu -q -x --timeout=300 -k "conv_stem or engine or smoke or checkpoint" > gpurun_out/pytest31.log 2>&1; echo "pytest rc=$?" > gpurun_out/rc31.txt
timeout 300 python benchmarks/resnet_breakdown.py --top 14 --out gpurun_out/resnet_breakdown2.json > gpurun_out/resnet_breakdown2.log 2>&1; echo "breakdown rc=$?" >> gpurun_out/rc31.txt
timeout 600 python bench.py > gpurun_out/bench31.json 2> gpurun_out/bench31.err; echo "bench rc=$?" >> gpurun_out/rc31.txt
cat gpurun_out/rc31.txt; tail -n 12 gpurun_out/pytest31.log; grep -v Warn gpurun_out/resnet_breakdown2.log | cut -c1-160 | head -26; cut -c1-400 gpurun_out/bench31.json
