set -x
python scripts/prof_conv_pair.py || exit 1
ncu --set full --clock-control none -k regex:"conv_tc3" -s 6 -c 3 -o /tmp/pair -f python scripts/prof_conv_pair.py > gpurun_out/ncu3.log 2>&1
ncu -i /tmp/pair.ncu-rep --page raw --csv > gpurun_out/r01_conv_pair_raw.csv 2>/dev/null
bash scripts/gpu_launches.sh
