#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -k "model or pipeline or driver or stress" 2>&1 | tail -3
timeout 900 python bench.py --no-x3 > gpurun_out/r2_e2e_v6_bench.json 2>gpurun_out/bench_err.log; tail -2 gpurun_out/bench_err.log
python -c "
import json; d=json.load(open('gpurun_out/r2_e2e_v6_bench.json'))
print({k:d[k] for k in ('value','ms_per_step')}, d['e2e']['value'], d.get('e2e_from_bam',{}).get('value'), d['roofline']['frac'], d['roofline']['ms_per_launch'], d['roofline']['traffic'], d.get('parity'))"
