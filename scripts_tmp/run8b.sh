TR8="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
echo "== multi-GPU pytest"; timeout 700 python -m pytest tests/test_multigpu_gpu.py -q -m gpu > gpurun_out/multigpu_8b.log 2>&1; tail -4 gpurun_out/multigpu_8b.log
echo "== allreduce + link bytes"; timeout 400 $TR8 --master-port 29605 benchmarks/allreduce_bench.py --link-bytes --out gpurun_out/allreduce_8gpu_v2.json > gpurun_out/allreduce_8gpu_v2.log 2>&1; python - <<'P'
import json
d=json.load(open('gpurun_out/allreduce_8gpu_v2.json'))
print(json.dumps(d.get('nvlink_counters_rank0'),indent=1)); print(d.get('nvls_inplace_64MB_us_by_blocks'))
for r in d['rows']:
    if r['bytes']>=(16<<20): print(r['bytes'], {k:v['us'] for k,v in r.items() if isinstance(v,dict)})
P
echo "== bert 8"; timeout 600 python scripts_tmp/bertjob8.py > gpurun_out/bert_8gpu.log 2>&1; tail -25 gpurun_out/bert_8gpu.log | cut -c1-300
