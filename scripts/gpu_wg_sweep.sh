python scripts/bench_wgrad_layer.py 2>&1 | head -6
OF_WG_NSTACK2=1 python scripts/bench_wgrad_layer.py 2>&1 | head -6
OF_WG_NSTACK2=1 python -m pytest tests/test_gpu_conv_tc.py -x -q -m gpu -k wgrad 2>&1 | tail -3
