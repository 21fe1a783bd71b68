#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q --timeout=120 > gpurun_out/pytest7_alone.log 2>&1; echo "alone rc=$?" >> gpurun_out/rc7.txt
timeout 900 python -m pytest tests -m gpu -q -x --timeout=300 > gpurun_out/pytest7_all.log 2>&1; echo "all rc=$?" >> gpurun_out/rc7.txt
cat gpurun_out/rc7.txt; tail -n 5 gpurun_out/pytest7_alone.log gpurun_out/pytest7_all.log
