#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_encoder_gpu.py -q -x > gpurun_out/r2_job4_enc.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_job4_enc.log
tail -25 gpurun_out/r2_job4_enc.log
DP_ENCODER=v4 timeout 600 python scripts/enc_stress.py 60 5000 > gpurun_out/r2_job4_stress.log 2>&1; echo "stress rc=$?" >> gpurun_out/r2_job4_stress.log
tail -5 gpurun_out/r2_job4_stress.log
for enc in v3 v4; do for nw in 2048 4096; do
  echo "== DP_ENCODER=$enc windows=$nw" >> gpurun_out/r2_job4_bench.log
  DP_ENCODER=$enc DP_ENC_TIMING=0 timeout 300 python bench.py --stage encode --steps 20 --warmup 3 --windows $nw --no-cpu-baseline >> gpurun_out/r2_job4_bench.log 2>&1
done; done
DP_ENCODER=v4 DP_ENC_TIMING=1 timeout 300 python bench.py --stage encode --steps 2 --warmup 3 --windows 4096 --no-cpu-baseline 2>&1 | grep "dp_encode_windows" | tail -3
python - <<'PY'
import json
for ln in open('gpurun_out/r2_job4_bench.log'):
    if ln.startswith('=='): print(ln.strip())
    if ln.startswith('{'):
        d=json.loads(ln); r=d['roofline']; print(' windows/s %.4g  ms/launch %.4f frac %.3f fallback %s' % (d['windows_per_s'], r['ms_per_launch'], r['frac'], r.get('fallback_windows_last_shard')))
PY
tail -3 gpurun_out/r2_job4_bench.log | cut -c1-300
