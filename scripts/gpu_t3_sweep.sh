python scripts/bench_conv_layer.py
echo "OF_T3_2CTA=1"
OF_T3_2CTA=1 python scripts/bench_conv_layer.py
OF_T3_2CTA=1 python -m pytest tests/test_gpu_conv_tc.py -x -q -m gpu 2>&1 | tail -2
