#!/bin/bash
# GPU job at the end of a round: parity tests, smoke, the two bench lines, the ncu launch list of the bench command and
# one `ncu --set full` capture of the dominant kernel (each ncu pass only after the same command exited 0 without ncu).
mkdir -p gpurun_out
T=${1:-v4}
timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -3
timeout 300 python __graft_entry__.py --smoke 2>&1 | tail -2
timeout 600 python bench.py > gpurun_out/bench_e2e_$T.json 2> gpurun_out/bench_e2e_$T.err; echo "bench rc=$?"
timeout 600 python bench.py --stage encode --steps 40 --warmup 3 --windows 4096 > gpurun_out/bench_enc_$T.json 2>> gpurun_out/bench_e2e_$T.err; echo "bench enc rc=$?"
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref_$T.json 2>> gpurun_out/bench_e2e_$T.err; echo "bench ref rc=$?"
CMD="python bench.py --steps 2 --warmup 3 --windows 1024 --no-cpu-baseline"
timeout 300 $CMD > gpurun_out/plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$T.csv $CMD > gpurun_out/ncu_launch.log 2>&1
echo "ncu launch list rc=$?"
timeout 300 $CMD > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gru_step2 -s 20 -c 2 -f -o gpurun_out/gru_full_$T $CMD > gpurun_out/ncu_gru.log 2>&1
echo "ncu full rc=$?"
ls -la gpurun_out | tail -8
