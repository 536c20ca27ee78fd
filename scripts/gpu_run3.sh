set -x
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > /dev/null 2>&1 || exit 1
ncu --set full --clock-control none -k regex:"cv7|colsum|copy_channels|sample_bwd|smallcout|act_bwd|zero_upsample|gemm_simt" -c 60 -o /tmp/misc_full -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu3.log 2>&1
echo rc $?
ncu -i /tmp/misc_full.ncu-rep --page raw --csv > gpurun_out/r01_misc_raw.csv 2>/dev/null
ls -la /tmp/misc_full.ncu-rep gpurun_out/
