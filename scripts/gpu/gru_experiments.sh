#!/bin/bash
# timing experiments on the GRU step kernels (results are WRONG under DP_GRU_DEBUG != 0; timing only)
for cfg in "1cta 6" "1cta 7" "1cta 8" "1cta 9" "1cta 5"; do
  set -- $cfg
  echo "== DP_GRU=$1 DP_GRU_DEBUG=$2"
  DP_GRU=$1 DP_GRU_DEBUG=$2 timeout 200 python bench.py --steps 3 --warmup 3 --no-cpu-baseline 2>&1 | python -c "
import sys,json
for ln in sys.stdin:
    if ln.startswith('{'):
        d=json.loads(ln); print('value',d['value'],'gru ms/launch',d['roofline']['ms_per_launch'],'TF',d['roofline']['achieved'])
"
done
