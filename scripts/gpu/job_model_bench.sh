#!/bin/bash
# GPU job: parity tests, smoke, bench (plain), then the ncu launch list of the same bench command.
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -15
timeout 300 python __graft_entry__.py --smoke 2>&1 | tail -3
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_e2e.json 2> gpurun_out/bench_e2e.err; tail -c 3000 gpurun_out/bench_e2e.json; tail -5 gpurun_out/bench_e2e.err
timeout 300 python bench.py --steps 2 --warmup 3 --windows 1024 --no-cpu-baseline > gpurun_out/plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 2 --warmup 3 --windows 1024 --no-cpu-baseline > gpurun_out/ncu.log 2>&1
tail -3 gpurun_out/ncu.log
