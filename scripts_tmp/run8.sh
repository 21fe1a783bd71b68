TR8="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
nvidia-smi topo -m > gpurun_out/topo8.txt 2>&1
echo "== multi-GPU pytest"; timeout 700 python -m pytest tests/test_multigpu_gpu.py -q -m gpu > gpurun_out/multigpu_8.log 2>&1; tail -4 gpurun_out/multigpu_8.log
echo "== bench 8"; timeout 400 $TR8 --master-port 29601 bench.py --gpus 8 --steps 40 --warmup 5 > gpurun_out/bench_8gpu.json 2> gpurun_out/bench_8gpu.err; cut -c1-600 gpurun_out/bench_8gpu.json
echo "== bench 8 comm-blocks 16"; timeout 300 $TR8 --master-port 29602 bench.py --gpus 8 --steps 40 --warmup 5 --comm-blocks 16 --no-via-launcher --launch-trials 0 > gpurun_out/bench_8gpu_cb16.json 2> gpurun_out/bench_8gpu_cb16.err; cut -c1-200 gpurun_out/bench_8gpu_cb16.json
echo "== diag no allreduce"; timeout 300 $TR8 --master-port 29603 bench.py --gpus 8 --steps 40 --warmup 5 --diag-no-allreduce --diag-step-times --no-via-launcher --launch-trials 0 > gpurun_out/diag8_noar_steps.json 2> gpurun_out/diag8_noar_steps.err; python -c "
import json;d=json.loads(open('gpurun_out/diag8_noar_steps.json').read().strip().splitlines()[-1]);print(d['ms_per_step'],d['ms_per_step_per_rank'],d['step_time_spread'],d['clocks'])"
echo "== 8 solo at once"; for i in 0 1 2 3 4 5 6 7; do CUDA_VISIBLE_DEVICES=$i timeout 300 python bench.py --steps 40 --warmup 5 --no-via-launcher --launch-trials 0 > gpurun_out/solo8_$i.json 2> gpurun_out/solo8_$i.err & done; wait; for i in 0 1 2 3 4 5 6 7; do python -c "
import json;d=json.loads(open('gpurun_out/solo8_$i.json').read().strip().splitlines()[-1]);print($i,d['value'],d['ms_per_step'],d['clocks']['sm_mhz'])"; done
echo "== gemm_allreduce"; timeout 400 $TR8 --master-port 29604 benchmarks/gemm_allreduce_bench.py > gpurun_out/gar_8gpu.json 2> gpurun_out/gar_8gpu.err; grep "tile128\|fused\|unfused nvls\|cublas\|^[0-9]" gpurun_out/gar_8gpu.err
echo "== allreduce"; timeout 400 $TR8 --master-port 29605 benchmarks/allreduce_bench.py --out gpurun_out/allreduce_8gpu.json > gpurun_out/allreduce_8gpu.log 2>&1; tail -3 gpurun_out/allreduce_8gpu.log | cut -c1-400
echo "== wide deep"; timeout 500 python benchmarks/wide_deep_job.py --ps 2 --workers 6 --out gpurun_out/wide_deep_8gpu.json > gpurun_out/wide_deep_8gpu.log 2>&1; grep "samples_per_s\|ms_per_step\|native_over\|error" gpurun_out/wide_deep_8gpu.log
