#!/bin/bash
# Round-end verification on one GPU: full GPU test suite, smoke(), bench.py (native + reference arm)
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x --timeout=400 > gpurun_out/pytest_final.log 2>&1; echo "pytest rc=$?" > gpurun_out/rc_final.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_final.log 2>&1; echo "smoke rc=$?" >> gpurun_out/rc_final.txt
timeout 600 python bench.py > gpurun_out/bench_final.json 2> gpurun_out/bench_final.err; echo "bench rc=$?" >> gpurun_out/rc_final.txt
timeout 120 python bench.py --impl reference > gpurun_out/bench_ref_final.json 2>/dev/null; echo "bench ref rc=$?" >> gpurun_out/rc_final.txt
cat gpurun_out/rc_final.txt; tail -n 4 gpurun_out/pytest_final.log; tail -n 2 gpurun_out/smoke_final.log; cut -c1-330 gpurun_out/bench_final.json; cat gpurun_out/bench_ref_final.json
