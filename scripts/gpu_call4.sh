#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x --timeout=300 > gpurun_out/pytest4.log 2>&1; echo "pytest rc=$?" >> gpurun_out/rc4.txt
timeout 300 python benchmarks/stem_probe.py > gpurun_out/stem_probe.json 2> gpurun_out/stem_probe.err; echo "stem rc=$?" >> gpurun_out/rc4.txt
timeout 300 python benchmarks/kernel_bench.py --out gpurun_out/kernel_bench4.json > gpurun_out/kernel_bench4.log 2>&1; echo "kb rc=$?" >> gpurun_out/rc4.txt
timeout 400 python bench.py --steps 20 --warmup 5 > gpurun_out/bench4.json 2> gpurun_out/bench4.err; echo "bench rc=$?" >> gpurun_out/rc4.txt
cat gpurun_out/rc4.txt; tail -n 3 gpurun_out/pytest4.log; cat gpurun_out/stem_probe.json gpurun_out/bench4.json
