#!/bin/bash
# round 2, call 1: full GPU suite (timed), default bench line, loss workload with both flow statistics
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/r2_smi.txt 2>&1
( time timeout 1500 python -m pytest tests -m gpu -x -q --durations=25 ) > gpurun_out/r2_pytest_1.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2_pytest_1.log
timeout 600 python bench.py > gpurun_out/r2_bench_1.json 2> gpurun_out/r2_bench_1.err
timeout 300 python bench.py --workload loss --flows iid --steps 200 --no-cpu-baseline > gpurun_out/r2_loss_iid_1.json 2>> gpurun_out/r2_bench_1.err
timeout 300 python bench.py --workload loss --flows smooth --steps 200 --no-cpu-baseline > gpurun_out/r2_loss_smooth_1.json 2>> gpurun_out/r2_bench_1.err
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_smoke_1.log 2>&1
tail -5 gpurun_out/r2_pytest_1.log; cat gpurun_out/r2_bench_1.json | head -c 3000; tail -3 gpurun_out/r2_smoke_1.log
