timeout 600 python -m pytest tests -m gpu -q --timeout 180 > gpurun_out/t_all.log 2>&1; tail -3 gpurun_out/t_all.log
for fl in smooth iid; do echo $fl; timeout 300 python bench.py --workload loss --no-cpu-baseline --steps 200 --flows $fl 2>&1 | grep -o '"ms_per_step": [0-9.]*'; done
OF_LF_DBG=5 timeout 300 python bench.py --workload loss --no-cpu-baseline --steps 200 --flows smooth 2>&1 | grep -o '"ms_per_step": [0-9.]*'
OF_LOSS_GENERIC=1 timeout 300 python bench.py --workload loss --no-cpu-baseline --steps 200 --flows smooth 2>&1 | grep -o '"ms_per_step": [0-9.]*'
