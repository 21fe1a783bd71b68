#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x --timeout=400 > gpurun_out/pytest30.log 2>&1; echo "pytest rc=$?" > gpurun_out/rc30.txt
timeout 600 python bench.py > gpurun_out/bench30.json 2> gpurun_out/bench30.err; echo "bench rc=$?" >> gpurun_out/rc30.txt
timeout 200 python benchmarks/gemm_once.py > gpurun_out/gemm_once.log 2>&1; echo "gemm_once rc=$?" >> gpurun_out/rc30.txt
if tail -n 1 gpurun_out/rc30.txt | grep -q "rc=0"; then
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_tn_kernel -s 2 -c 1 -o gpurun_out/gemm_2cta python benchmarks/gemm_once.py > gpurun_out/gemm_ncu.log 2>&1; echo "ncu rc=$?" >> gpurun_out/rc30.txt
ncu -i gpurun_out/gemm_2cta.ncu-rep --page raw --csv > gpurun_out/gemm_2cta_raw.csv 2>/dev/null
fi
cat gpurun_out/rc30.txt; tail -n 6 gpurun_out/pytest30.log; cat gpurun_out/bench30.json; tail -n 3 gpurun_out/bench30.err
