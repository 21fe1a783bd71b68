#!/bin/bash
# ncu evidence: launch list of one training step, full captures of the BN kernels and the GEMM
set -x
mkdir -p gpurun_out
timeout 400 python bench.py --steps 10 --warmup 3 --impl stock --stock-mode autocast > gpurun_out/bench_stock_amp.json 2> gpurun_out/bench_stock_amp.err; echo "stock amp rc=$?" >> gpurun_out/rc3.txt
CMD="python bench.py --no-graph --steps 1 --warmup 3"
timeout 300 $CMD > gpurun_out/plain_step.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 2400 -c 900 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "launches rc=$?" >> gpurun_out/rc3.txt
timeout 900 ncu --set full --clock-control none --import-source on -k regex:bn_ -s 40 -c 8 -f -o gpurun_out/prof_bn $CMD > gpurun_out/ncu_bn.log 2>&1
echo "ncu bn rc=$?" >> gpurun_out/rc3.txt
timeout 120 python benchmarks/gemm_once.py > gpurun_out/gemm_once.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_tn -s 2 -c 1 -f -o gpurun_out/prof_gemm python benchmarks/gemm_once.py > gpurun_out/ncu_gemm.log 2>&1
echo "ncu gemm rc=$?" >> gpurun_out/rc3.txt
cat gpurun_out/rc3.txt; cat gpurun_out/bench_stock_amp.json; ls -la gpurun_out | head -40
