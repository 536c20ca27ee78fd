set -x
timeout 600 python -m pytest tests -m gpu -q --timeout 120 > gpurun_out/t_all.log 2>&1; tail -5 gpurun_out/t_all.log
timeout 300 python bench.py --workload loss --no-cpu-baseline > gpurun_out/bench_loss.log 2>&1; tail -3 gpurun_out/bench_loss.log
timeout 300 python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/plain.log 2>&1 && timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches_train.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_train.log 2>&1
tail -3 gpurun_out/ncu_train.log; wc -l gpurun_out/launches_train.csv
