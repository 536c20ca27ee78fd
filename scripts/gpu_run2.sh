timeout 900 python -m pytest tests/test_gpu_conv_tc.py tests/test_gpu_net.py -m gpu -q --timeout 180 > gpurun_out/t_tc.log 2>&1; tail -3 gpurun_out/t_tc.log
timeout 600 python bench.py --no-cpu-baseline 2>&1 | tail -1 | grep -o '"ms_per_step": [0-9.]*'
