#!/bin/bash
# ncu of the encoder kernels of one dp_encode_windows call (4096 windows): launch list, then one `--set full` capture of
# prep_reads / base_codes / encode_windows_v3 (each ncu pass only after the same command exited 0 without ncu)
mkdir -p gpurun_out
T=${1:-v6}
CMD="python bench.py --stage encode --steps 2 --warmup 3 --windows 4096 --no-cpu-baseline"
timeout 300 $CMD > gpurun_out/plain_enc.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file gpurun_out/launches_enc_$T.csv $CMD > gpurun_out/ncu_enc.log 2>&1
echo "launch list rc=$?"
timeout 300 $CMD > /dev/null 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"prep_reads|base_codes|encode_windows_v3" -s 12 -c 3 -f -o gpurun_out/enc_full_$T $CMD > gpurun_out/ncu_enc_full.log 2>&1
echo "full rc=$?"; tail -2 gpurun_out/ncu_enc_full.log
