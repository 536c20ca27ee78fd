timeout 300 python bench.py --workload loss --no-cpu-baseline --steps 8 --no-graph > gpurun_out/plain_loss.log 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:loss_fused -s 4 -c 1 -o gpurun_out/prof_loss -f python bench.py --workload loss --no-cpu-baseline --steps 8 --no-graph > gpurun_out/ncu_loss.log 2>&1
tail -2 gpurun_out/ncu_loss.log
