python scripts/bench_wgrad_layer.py
OF_WG_WAVES=1 python scripts/bench_wgrad_layer.py 2>&1 | head -6
OF_WG_WAVES=3 python scripts/bench_wgrad_layer.py 2>&1 | head -6
