python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/b1.json 2>gpurun_out/b1.err
OF_CV_NO_TC=1 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/b2.json 2>gpurun_out/b2.err
python scripts/bench_conv_layer.py > gpurun_out/b3.txt 2>&1
