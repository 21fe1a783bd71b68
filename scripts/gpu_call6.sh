#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x --timeout=300 > gpurun_out/pytest6.log 2>&1; echo "pytest rc=$?" >> gpurun_out/rc6.txt
timeout 400 python bench.py --steps 20 --warmup 5 > gpurun_out/bench6.json 2> gpurun_out/bench6.err; echo "bench rc=$?" >> gpurun_out/rc6.txt
timeout 400 python bench.py --steps 20 --warmup 5 --no-s2d-stem > gpurun_out/bench6_nos2d.json 2> gpurun_out/bench6_nos2d.err; echo "bench nos2d rc=$?" >> gpurun_out/rc6.txt
cat gpurun_out/rc6.txt; tail -n 3 gpurun_out/pytest6.log; tail -n 3 gpurun_out/bench6.err; cat gpurun_out/bench6.json gpurun_out/bench6_nos2d.json
