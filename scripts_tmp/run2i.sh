TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 150 $TR --master-port 29581 benchmarks/ps_bench.py --restrict --no-nccl-first --out gpurun_out/ps_bench_2gpu_v5.json > gpurun_out/ps_bench_2gpu_v5.log 2>&1; python -c "
import json;d=json.load(open('gpurun_out/ps_bench_2gpu_v5.json'));print(json.dumps(d['remote_gather_fresh_ids_vs_idle_gap'],indent=1))"; tail -3 gpurun_out/ps_bench_2gpu_v5.log | cut -c1-300
