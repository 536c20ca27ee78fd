set -x
timeout 600 python -m pytest tests/test_gpu_ops.py tests/test_golden.py -m gpu -q --timeout 180 -k "loss" > gpurun_out/t_loss.log 2>&1; tail -8 gpurun_out/t_loss.log
timeout 300 python bench.py --workload loss --no-cpu-baseline --steps 200 > gpurun_out/bench_loss.log 2>&1; tail -3 gpurun_out/bench_loss.log
