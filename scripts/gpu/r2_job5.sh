#!/bin/bash
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -q > gpurun_out/r2_job5_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_job5_tests.log
grep -E "passed|failed|FAILED|ERROR|rc=" gpurun_out/r2_job5_tests.log | tail -15
timeout 900 python scripts/bgzf_bench.py --kb 6000 > gpurun_out/r2_bgzf_v4_bench_6000kb.json 2> gpurun_out/r2_bgzf_v4.err; tail -2 gpurun_out/r2_bgzf_v4.err
cat gpurun_out/r2_bgzf_v4_bench_6000kb.json | head -40
DP_BAM_SCAN=serial timeout 900 python scripts/bgzf_bench.py --kb 1500 > gpurun_out/r2_bgzf_v4_bench_serial_scan.json 2>/dev/null
timeout 900 python scripts/bgzf_bench.py --kb 1500 > gpurun_out/r2_bgzf_v4_bench_1500kb.json 2>/dev/null
python - <<'PY'
import json
for f in ['gpurun_out/r2_bgzf_v4_bench_serial_scan.json','gpurun_out/r2_bgzf_v4_bench_1500kb.json']:
    try:
        d=json.load(open(f)); print(f, {k:v for k,v in d.items() if 'ms' in k or 'gbs' in k.lower() or 'GB' in k})
    except Exception as e: print(f, e)
PY
for nw in 2048 4096; do
  echo "== windows=$nw" >> gpurun_out/r2_job5_bench.log
  DP_ENC_TIMING=0 timeout 300 python bench.py --stage encode --steps 20 --warmup 3 --windows $nw --no-cpu-baseline >> gpurun_out/r2_job5_bench.log 2>&1
done
DP_ENC_TIMING=1 timeout 300 python bench.py --stage encode --steps 2 --warmup 3 --windows 4096 --no-cpu-baseline 2>&1 | grep "dp_encode_windows" | tail -2
python - <<'PY'
import json
for ln in open('gpurun_out/r2_job5_bench.log'):
    if ln.startswith('=='): print(ln.strip())
    if ln.startswith('{'):
        d=json.loads(ln); r=d['roofline']; print(' windows/s %.4g  ms/launch %.4f frac %.3f fallback %s' % (d['windows_per_s'], r['ms_per_launch'], r['frac'], r.get('fallback_windows_last_shard')))
PY
tail -2 gpurun_out/r2_job5_bench.log | cut -c1-400
