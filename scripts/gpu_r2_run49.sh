#!/bin/bash
# round 2: conv_mode="bf16x3" (tensor-core parity mode): tests, then its step time beside bf16 / fp32
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_b6_gpu_bf16x3.py -m gpu -x -q 2>&1 | tail -15
