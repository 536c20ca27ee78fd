timeout 900 python -m pytest tests -m gpu -q --timeout 180 > gpurun_out/t_all.log 2>&1; tail -25 gpurun_out/t_all.log
for fl in smooth iid; do echo $fl; timeout 300 python bench.py --workload loss --no-cpu-baseline --steps 200 --flows $fl 2>&1 | grep -o '"ms_per_step": [0-9.]*'; done
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/bench_train.log 2>&1; tail -2 gpurun_out/bench_train.log
