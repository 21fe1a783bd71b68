#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 300 python benchmarks/all_kernels_once.py > gpurun_out/all_once.log 2>&1; echo "plain rc=$?" > gpurun_out/rc33.txt
if grep -q "rc=0" gpurun_out/rc33.txt; then
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"stem|u8_to|maxpool|sgd_momentum|adamw|multi_pack|zero_kernel|gemm_tn_kernel|transpose|xent|ps_|hist|allreduce|one_shot|two_shot|gemm_ar|gar" -o gpurun_out/all_kernels python benchmarks/all_kernels_once.py > gpurun_out/all_ncu.log 2>&1; echo "ncu rc=$?" >> gpurun_out/rc33.txt
ncu -i gpurun_out/all_kernels.ncu-rep --page raw --csv > gpurun_out/all_kernels_raw.csv 2>/dev/null
fi
cat gpurun_out/rc33.txt; tail -n 3 gpurun_out/all_once.log; tail -n 3 gpurun_out/all_ncu.log; wc -l gpurun_out/all_kernels_raw.csv
