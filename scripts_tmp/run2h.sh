TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 150 $TR --master-port 29581 benchmarks/ps_bench.py --restrict --no-nccl-first --out gpurun_out/ps_bench_2gpu_v3.json > gpurun_out/ps_bench_2gpu_v3.log 2>&1; tail -1 gpurun_out/ps_bench_2gpu_v3.log | cut -c1-1800
timeout 150 $TR --master-port 29582 benchmarks/ps_bench.py --no-nccl-first --out gpurun_out/ps_bench_2gpu_v4.json > gpurun_out/ps_bench_2gpu_v4.log 2>&1; tail -1 gpurun_out/ps_bench_2gpu_v4.log | cut -c1-1800
