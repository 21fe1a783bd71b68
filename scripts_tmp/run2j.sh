TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
for cfg in "256 10000016" "256 10625000" "2097152 10625000"; do set -- $cfg
echo "== align $1 rows $2"
TFOP_COMM_ALIGN=$1 timeout 150 $TR --master-port 29591 benchmarks/ps_bench.py --restrict --no-nccl-first --rows $2 --out gpurun_out/ps_bench_align.json > gpurun_out/ps_bench_align.log 2>&1; python -c "
import json;d=json.load(open('gpurun_out/ps_bench_align.json'));print(d['remote_owner_gpu_idle'], d['local']['gather_us'])"
done
timeout 300 python benchmarks/wide_deep_job.py --ps 1 --workers 1 --out gpurun_out/wide_deep_2gpu_v4.json > gpurun_out/wide_deep_2gpu_v4.log 2>&1; grep "samples_per_s\|ms_per_step\|native_over\|error" gpurun_out/wide_deep_2gpu_v4.log
