#!/bin/bash
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --windows 1024 --no-cpu-baseline --no-x3 --no-parity --no-bam"
timeout 600 $CMD > gpurun_out/r2_prof_cmd.json 2>/dev/null || exit 1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_e2e_v6_launches.csv $CMD > /dev/null 2>&1
timeout 900 ncu --set full --clock-control none -k regex:gru_step2 -s 20 -c 2 -o gpurun_out/r2_gru_step2_v6 -f $CMD > /dev/null 2>&1
ncu -i gpurun_out/r2_gru_step2_v6.ncu-rep --page raw --csv > gpurun_out/r2_gru_step2_v6_raw.csv 2>/dev/null
ls -la gpurun_out | tail -5
