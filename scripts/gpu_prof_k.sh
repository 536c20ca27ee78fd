# ncu --set full on the launches of one kernel (regex $1), skipping $2 launches, capturing $3: summary -> gpurun_out/$4_raw.csv
set -x
ncu --set full --clock-control none -k regex:"$1" -s $2 -c $3 -o /tmp/k -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_k.log 2>&1
ncu -i /tmp/k.ncu-rep --page raw --csv > gpurun_out/$4_raw.csv 2>/dev/null
ls -la /tmp/k.ncu-rep gpurun_out/$4_raw.csv
