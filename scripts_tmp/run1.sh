timeout 1500 python -m pytest tests/ -q -m gpu -x > gpurun_out/gpu_all_final.log 2>&1; tail -4 gpurun_out/gpu_all_final.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke_final.log 2>&1; tail -2 gpurun_out/smoke_final.log
timeout 600 python bench.py > gpurun_out/bench_1gpu_final.json 2> gpurun_out/bench_1gpu_final.err; cut -c1-700 gpurun_out/bench_1gpu_final.json
timeout 300 python examples/bert_large_dp/train.py --steps 20 --warmup 4 > gpurun_out/bert_1gpu_final.log 2>&1; tail -1 gpurun_out/bert_1gpu_final.log
timeout 120 python bench.py --impl reference; 
