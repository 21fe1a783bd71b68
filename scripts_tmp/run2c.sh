TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 400 python -m pytest tests/test_multigpu_gpu.py -q -m gpu -k "collective or ps_worker" > gpurun_out/mg_kernels_2c.log 2>&1; tail -5 gpurun_out/mg_kernels_2c.log
timeout 400 $TR --master-port 29531 benchmarks/gemm_allreduce_bench.py > gpurun_out/gar_2gpu_v2.json 2> gpurun_out/gar_2gpu_v2.err; grep -v "^ *$" gpurun_out/gar_2gpu_v2.err | tail -45
timeout 600 python benchmarks/wide_deep_job.py --ps 1 --workers 1 --out gpurun_out/wide_deep_2gpu.json > gpurun_out/wide_deep_2gpu.log 2>&1; tail -30 gpurun_out/wide_deep_2gpu.log
