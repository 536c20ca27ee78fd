set -x
python scripts/prof_wgrad.py || exit 1
ncu --set full --clock-control none -k regex:"conv_tc_wgrad" -o /tmp/wg -f python scripts/prof_wgrad.py > gpurun_out/ncu_wg.log 2>&1
ncu -i /tmp/wg.ncu-rep --page raw --csv > gpurun_out/wgrad_raw.csv 2>/dev/null
ls -la /tmp/wg.ncu-rep gpurun_out/wgrad_raw.csv
