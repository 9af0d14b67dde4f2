#!/bin/bash
# ncu --set full capture of the GRU step kernel (and the folded GEMM) on a short bench run
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --windows 512 --no-cpu-baseline"
timeout 300 $CMD > gpurun_out/plain_ncu.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"gru_step|gemm_bf16" -s 24 -c 4 -f -o gpurun_out/prof_model $CMD > gpurun_out/ncu_full.log 2>&1
tail -5 gpurun_out/ncu_full.log
ls -la gpurun_out/
