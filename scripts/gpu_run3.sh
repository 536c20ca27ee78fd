timeout 300 python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/plain.log 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on -k regex:conv_tc3 -s 214 -c 2 -o gpurun_out/prof_tc3 -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_tc3.log 2>&1
tail -2 gpurun_out/ncu_tc3.log
