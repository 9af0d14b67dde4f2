#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -k "bgzf or pipeline" 2>&1 | tail -2
timeout 900 python scripts/bgzf_bench.py --kb 6000 > gpurun_out/r2_bgzf_v6_bench_6000kb.json 2>gpurun_out/r2_bgzf_v6.err; tail -2 gpurun_out/r2_bgzf_v6.err
python -c "
import json; d=json.load(open('gpurun_out/r2_bgzf_v6_bench_6000kb.json')); print({k:(round(v,3) if isinstance(v,float) else v) for k,v in d.items() if k!='note' and k!='workload'})"
