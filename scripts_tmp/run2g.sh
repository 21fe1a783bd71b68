TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 60 python -m pytest tests/test_kernels_gpu.py -q -m gpu -k "serve" > gpurun_out/t_ps5.log 2>&1; tail -5 gpurun_out/t_ps5.log
timeout 150 $TR --master-port 29571 benchmarks/ps_bench.py --restrict --out gpurun_out/ps_bench_2gpu_v2.json > gpurun_out/ps_bench_2gpu_v2.log 2>&1; tail -1 gpurun_out/ps_bench_2gpu_v2.log | cut -c1-1500
timeout 300 python benchmarks/wide_deep_job.py --ps 1 --workers 1 --out gpurun_out/wide_deep_2gpu_v3.json > gpurun_out/wide_deep_2gpu_v3.log 2>&1; grep "samples_per_s\|ms_per_step\|native_over\|error" gpurun_out/wide_deep_2gpu_v3.log
