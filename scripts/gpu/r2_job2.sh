#!/bin/bash
# round 2, job 2: full GPU suite with the new tests, the new bench line, staggered-GRU ring sweep
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -s > gpurun_out/r2_job2_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_job2_tests.log
timeout 600 python bench.py > gpurun_out/r2_bench_v1.json 2> gpurun_out/r2_bench_v1.err
for cfg in 0 1 2 3 4 5; do
  echo "== stag cfg=$cfg" >> gpurun_out/r2_job2_stag.log
  DP_GRU=stag DP_GRUS_CFG=$cfg timeout 300 python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-x3 --no-parity >> gpurun_out/r2_job2_stag.log 2>&1
done
for dbg in "1 0" "0 2"; do
  set -- $dbg
  echo "== stag cfg=0 DP_GRU_DEBUG=$1 DP_GRU_DEBUG2=$2" >> gpurun_out/r2_job2_stag.log
  DP_GRU=stag DP_GRUS_CFG=0 DP_GRU_DEBUG=$1 DP_GRU_DEBUG2=$2 timeout 300 python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-x3 --no-parity >> gpurun_out/r2_job2_stag.log 2>&1
done
grep -E "passed|failed|error|parity|test.py loop|entries moved|rc=" gpurun_out/r2_job2_tests.log | tail -30
python - <<'PY'
import json
for f in ['gpurun_out/r2_bench_v1.json','gpurun_out/r2_job2_stag.log']:
    for ln in open(f):
        if ln.startswith('=='): print(ln.strip())
        if ln.startswith('{'):
            d=json.loads(ln); r=d['roofline']; print(' value %.4g e2e %.4g gru ms/launch %.4f frac %.3f sm %s' % (d['value'], d['e2e']['value'], r['ms_per_launch'], r['frac'], d['clocks']['sm_mhz']))
            for k in d:
                if k.startswith('value_') or k.startswith('e2e_') or k.startswith('parity'): print('  ',k,d[k])
            if 'cpu_baseline' in d: print('  cpu', d['cpu_baseline']['value'], d['cpu_baseline']['legs'])
            if 'roofline_encoder' in d: print('  enc', d['roofline_encoder']['frac'], d['roofline_encoder']['ms_per_launch'])
PY
tail -5 gpurun_out/r2_bench_v1.err
