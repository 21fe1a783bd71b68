TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 300 python -m pytest tests/test_multigpu_gpu.py -q -m gpu -k collective > gpurun_out/mg_kernels_2.log 2>&1; tail -5 gpurun_out/mg_kernels_2.log
timeout 400 $TR --master-port 29531 benchmarks/gemm_allreduce_bench.py > gpurun_out/gar_2gpu.json 2> gpurun_out/gar_2gpu.err; tail -40 gpurun_out/gar_2gpu.err
timeout 300 $TR --master-port 29532 bench.py --gpus 2 --steps 40 --warmup 5 --diag-no-allreduce --diag-step-times --no-via-launcher --launch-trials 0 > gpurun_out/diag2_noar_steps.json 2> gpurun_out/diag2_noar_steps.err
timeout 300 $TR --master-port 29533 bench.py --gpus 2 --steps 40 --warmup 5 --diag-step-times --no-via-launcher --launch-trials 0 > gpurun_out/diag2_ar_steps.json 2> gpurun_out/diag2_ar_steps.err
for f in diag2_noar_steps diag2_ar_steps; do echo "== $f"; python - <<P
import json
try:
    d=json.loads(open("gpurun_out/$f.json").read().strip().splitlines()[-1])
    print(d.get("value"), d.get("ms_per_step"), d.get("ms_per_step_per_rank"), d.get("clocks"), d.get("step_time_spread"))
except Exception as e:
    print("ERR", e)
P
done
