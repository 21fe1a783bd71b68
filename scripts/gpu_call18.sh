#!/bin/bash
set -x
mkdir -p gpurun_out
CMD="python examples/bert_large_dp/train.py --steps 1 --warmup 2 --no-graph"
timeout 200 $CMD > gpurun_out/bert1_plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 2200 -c 2600 --csv --log-file gpurun_out/bert_launches.csv $CMD > gpurun_out/bert_ncu.log 2>&1
echo "bert launches rc=$?" >> gpurun_out/rc18.txt
wc -l gpurun_out/bert_launches.csv
