python scripts/bench_conv_layer.py 2>&1
python -m pytest tests/test_gpu_conv_tc.py -x -q -m gpu 2>&1 | tail -2
