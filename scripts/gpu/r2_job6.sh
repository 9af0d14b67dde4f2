#!/bin/bash
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -q -k "bgzf or windows or stress or pipeline" > gpurun_out/r2_job6_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_job6_tests.log
grep -E "passed|failed|FAILED|ERROR|rc=" gpurun_out/r2_job6_tests.log | tail -15
timeout 900 python scripts/bgzf_bench.py --kb 6000 > gpurun_out/r2_bgzf_v5_bench_6000kb.json 2> gpurun_out/r2_bgzf_v5.err; tail -2 gpurun_out/r2_bgzf_v5.err
cat gpurun_out/r2_bgzf_v5_bench_6000kb.json | cut -c1-700
timeout 900 python scripts/bgzf_bench.py --kb 1500 > gpurun_out/r2_bgzf_v5_bench_1500kb.json 2>/dev/null; cat gpurun_out/r2_bgzf_v5_bench_1500kb.json | cut -c1-700
timeout 2400 bash scripts/sanitize.sh > gpurun_out/r2_job6_sanitize.out 2>&1
cat gpurun_out/sanitize_summary.txt
