python scripts/bench_wgrad_layer.py 2>&1 | head -6
echo "stages 3, 2 waves (2 CTAs/SM where TMEM allows)"
OF_WG_STAGES=3 OF_WG_WAVES=2 python scripts/bench_wgrad_layer.py 2>&1 | head -6
echo "stages 2, 2 waves"
OF_WG_STAGES=2 OF_WG_WAVES=2 python scripts/bench_wgrad_layer.py 2>&1 | head -6
echo "stages 3, 1 wave"
OF_WG_STAGES=3 python scripts/bench_wgrad_layer.py 2>&1 | head -6
