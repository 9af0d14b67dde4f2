#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -k "pipeline or bgzf" 2>&1 | tail -1
timeout 900 python bench.py --no-x3 --no-parity > gpurun_out/r2_e2e_v5_bench.json 2>gpurun_out/bench_err.log; tail -2 gpurun_out/bench_err.log
python -c "
import json; d=json.load(open('gpurun_out/r2_e2e_v5_bench.json'))
print({k:d[k] for k in ('value','ms_per_step')}, d['e2e'], d.get('e2e_from_bam'))"
timeout 900 python scripts/config5_stream.py --chunks 6 --distinct 2 > gpurun_out/r2_config5_v5_1gpu.json 2>/dev/null
python -c "
import json; d=json.load(open('gpurun_out/r2_config5_v5_1gpu.json'))
print({k:d[k] for k in ('positions_per_s','seconds','ingest_ms_per_chunk_side_stream','compute_ms_per_chunk')})"
