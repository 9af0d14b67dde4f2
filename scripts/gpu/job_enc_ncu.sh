#!/bin/bash
mkdir -p gpurun_out
CMD="python bench.py --stage encode --steps 2 --warmup 3 --windows 4096 --no-cpu-baseline"
timeout 300 $CMD > gpurun_out/plain_enc.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches_enc.csv $CMD > gpurun_out/ncu_enc.log 2>&1
grep -E "^\"[0-9]" gpurun_out/launches_enc.csv | awk -F'","' '{print $5, $NF}' | tail -12
timeout 300 $CMD > /dev/null 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:encode_windows_v2 -s 3 -c 1 -f -o gpurun_out/prof_enc_v2 $CMD > gpurun_out/ncu_enc_full.log 2>&1
tail -2 gpurun_out/ncu_enc_full.log
