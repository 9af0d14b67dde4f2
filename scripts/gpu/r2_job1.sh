#!/bin/bash
# round 2, first GPU job: state of the suite, error attribution, GRU kernel variants under the bench
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,power.limit --format=csv > gpurun_out/r2_job1_gpu.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2_job1_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_job1_tests.log
timeout 900 python scripts/model_error_attribution.py --out gpurun_out/r2_model_error_attribution.json > gpurun_out/r2_job1_attr.log 2>&1
for cfg in "pair bf16" "pair fp16" "stag fp16" "wide fp16"; do
  set -- $cfg
  echo "== DP_GRU=$1 precision=$2" >> gpurun_out/r2_job1_bench.log
  DP_GRU=$1 timeout 300 python bench.py --steps 6 --warmup 3 --no-cpu-baseline --precision $2 >> gpurun_out/r2_job1_bench.log 2>&1
done
tail -5 gpurun_out/r2_job1_tests.log; cat gpurun_out/r2_job1_attr.log | tail -30
python - <<'PY'
import json
for ln in open('gpurun_out/r2_job1_bench.log'):
    if ln.startswith('=='): print(ln.strip())
    if ln.startswith('{'):
        d=json.loads(ln); r=d['roofline']; print(' value %.3g e2e %.3g gru ms/launch %.4f frac %.3f clocks %s' % (d['value'], d['e2e']['value'], r['ms_per_launch'], r['frac'], d['clocks']))
PY
