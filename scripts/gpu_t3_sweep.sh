for r in "" "2,8" "4,3" "3,5"; do echo "OF_T3_RINGS=$r"; OF_T3_RINGS=$r timeout 120 python scripts/bench_conv_layer.py 2>&1 | head -3; done
