TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 120 $TR --master-port 29561 benchmarks/ps_bench.py --restrict --out gpurun_out/ps_bench_2gpu_restrict.json > gpurun_out/ps_bench_2gpu_restrict.log 2>&1; tail -1 gpurun_out/ps_bench_2gpu_restrict.log | cut -c1-900
timeout 60 python -m pytest tests/test_kernels_gpu.py -q -m gpu -k "merges" > gpurun_out/t_ps3.log 2>&1; tail -15 gpurun_out/t_ps3.log
timeout 60 python -m pytest tests/test_kernels_gpu.py -q -m gpu -k "serve" > gpurun_out/t_ps4.log 2>&1; tail -15 gpurun_out/t_ps4.log
