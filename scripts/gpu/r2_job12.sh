#!/bin/bash
mkdir -p gpurun_out
timeout 900 python scripts/bgzf_bench.py --kb 12000 --sweep 2>/dev/null | tee gpurun_out/r2_inflate_v8_sweep.json
timeout 900 ncu --set full --import-source on --clock-control none -k regex:bgzf_inflate -c 1 -o gpurun_out/r2_inflate_v8 -f python scripts/bgzf_bench.py --kb 6000 --sweep > /dev/null 2>&1
ls -la gpurun_out/*.ncu-rep
