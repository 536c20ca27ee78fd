set -x
python -m pytest tests -x -q -m gpu 2>&1 | tail -4
python bench.py --steps 10 --warmup 3 --no-cpu-baseline 2>&1 | tail -1 | cut -c1-300
OF_CV_NO_TC=1 python bench.py --steps 10 --warmup 3 --no-cpu-baseline 2>&1 | tail -1 | cut -c1-300
