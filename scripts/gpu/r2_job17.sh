#!/bin/bash
mkdir -p gpurun_out
timeout 900 ncu --set full --import-source on --clock-control none -k regex:bgzf_inflate -c 1 -o gpurun_out/r2_inflate_v9_full -f python scripts/bgzf_bench.py --kb 6000 --reps 1 > /dev/null 2>&1
ls -la gpurun_out/*.ncu-rep
