TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
$TR --master-port 29521 bench.py --gpus 2 --steps 40 --warmup 5 --diag-no-allreduce --no-via-launcher --launch-trials 0 > gpurun_out/diag2_noar.json 2> gpurun_out/diag2_noar.err
$TR --master-port 29522 bench.py --gpus 2 --steps 40 --warmup 5 > gpurun_out/bench_2gpu_v2.json 2> gpurun_out/bench_2gpu_v2.err
$TR --master-port 29523 bench.py --gpus 2 --steps 40 --warmup 5 --comm-blocks 16 --no-via-launcher --launch-trials 0 > gpurun_out/diag2_cb16.json 2> gpurun_out/diag2_cb16.err
$TR --master-port 29524 bench.py --gpus 2 --steps 40 --warmup 5 --no-nvls --no-via-launcher --launch-trials 0 > gpurun_out/diag2_p2p.json 2> gpurun_out/diag2_p2p.err
CUDA_VISIBLE_DEVICES=0 python bench.py --steps 40 --warmup 5 --no-via-launcher --launch-trials 0 > gpurun_out/diag2_solo0.json 2> gpurun_out/diag2_solo0.err &
P0=$!
CUDA_VISIBLE_DEVICES=1 python bench.py --steps 40 --warmup 5 --no-via-launcher --launch-trials 0 > gpurun_out/diag2_solo1.json 2> gpurun_out/diag2_solo1.err &
P1=$!
wait $P0 $P1
CUDA_VISIBLE_DEVICES=0 python bench.py --steps 40 --warmup 5 --no-via-launcher --launch-trials 0 > gpurun_out/diag2_alone0.json 2> gpurun_out/diag2_alone0.err
for f in diag2_noar bench_2gpu_v2 diag2_cb16 diag2_p2p diag2_solo0 diag2_solo1 diag2_alone0; do echo "== $f"; python - <<P
import json
try:
    d=json.loads(open("gpurun_out/$f.json").read().strip().splitlines()[-1])
    print(d.get("value"), d.get("ms_per_step"), d.get("ms_per_step_per_rank"), d.get("clocks"), d.get("via_launcher"))
except Exception as e:
    print("ERR", e)
P
done
