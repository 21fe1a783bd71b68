TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29551 benchmarks/ps_bench.py --out gpurun_out/ps_bench_2gpu.json > gpurun_out/ps_bench_2gpu.log 2>&1; tail -3 gpurun_out/ps_bench_2gpu.log
timeout 300 python -m pytest tests/test_multigpu_gpu.py -q -m gpu -k "collective" > gpurun_out/mg_kernels_2d.log 2>&1; tail -3 gpurun_out/mg_kernels_2d.log
timeout 400 $TR --master-port 29552 benchmarks/gemm_allreduce_bench.py > gpurun_out/gar_2gpu_v3.json 2> gpurun_out/gar_2gpu_v3.err; grep "tile128\|fused nvls\|unfused nvls\|^[0-9]" gpurun_out/gar_2gpu_v3.err
