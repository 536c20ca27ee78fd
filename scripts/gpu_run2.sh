timeout 600 python -m pytest tests -x -q -m gpu 2>&1 | tail -6
python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/b1.json 2>gpurun_out/b1.err
OF_NO_PACK_CACHE=1 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/b2.json 2>gpurun_out/b2.err
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
