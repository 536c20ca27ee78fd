#!/bin/bash
# round 2, call 2: full GPU suite without -x (all failures at once), timed
mkdir -p gpurun_out
( time timeout 1500 python -m pytest tests -m gpu -q --durations=15 ) > gpurun_out/r2_pytest_2.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2_pytest_2.log
tail -40 gpurun_out/r2_pytest_2.log
