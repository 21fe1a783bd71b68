timeout 300 python -m pytest tests/test_kernels_gpu.py -q -m gpu -k "ps_" > gpurun_out/t_ps2.log 2>&1; tail -3 gpurun_out/t_ps2.log
timeout 600 python benchmarks/wide_deep_job.py --ps 1 --workers 1 --out gpurun_out/wide_deep_2gpu_v2.json > gpurun_out/wide_deep_2gpu_v2.log 2>&1; grep "samples_per_s\|ms_per_step\|native_over\|error" gpurun_out/wide_deep_2gpu_v2.log
timeout 300 python -m pytest tests/test_multigpu_gpu.py -q -m gpu -k "ps_worker" > gpurun_out/mg_ps.log 2>&1; tail -3 gpurun_out/mg_ps.log
