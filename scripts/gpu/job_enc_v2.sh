#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_encoder_gpu.py -x -q -m gpu 2>&1 | tail -15
timeout 300 python bench.py --stage encode --steps 5 --warmup 3 --windows 4096 --no-cpu-baseline 2>&1 | tail -c 1500
DP_ENCODER=v1 timeout 300 python bench.py --stage encode --steps 5 --warmup 3 --windows 4096 --no-cpu-baseline 2>&1 | tail -c 600
