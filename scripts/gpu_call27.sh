#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 200 python benchmarks/ln_once.py 10 > gpurun_out/ln_once.log 2>&1; echo "ln_once rc=$?" > gpurun_out/rc27.txt
if grep -q "rc=0" gpurun_out/rc27.txt; then
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"ln_bwd|ln_fwd|bias_act|bias_gelu|xent" -c 12 -o gpurun_out/ln_kernels python benchmarks/ln_once.py 1 > gpurun_out/ln_ncu.log 2>&1; echo "ncu rc=$?" >> gpurun_out/rc27.txt
ncu -i gpurun_out/ln_kernels.ncu-rep --page raw --csv > gpurun_out/ln_kernels_raw.csv 2>/dev/null
fi
cat gpurun_out/rc27.txt; tail -n 2 gpurun_out/ln_once.log; ls -la gpurun_out/ln_kernels* 
