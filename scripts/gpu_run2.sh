set -x
python -m pytest tests/test_gpu_conv_tc.py -x -q -m gpu 2>&1 | tail -5
for v in "" "OF_T3_NO_COAL=1" "OF_T3_DBG=1" "OF_T3_RINGS=2,6" "OF_T3_RINGS=3,3"; do echo "== $v"; env $v python scripts/bench_conv_layer.py; done
python bench.py --steps 10 --warmup 3 2>&1 | tail -2
