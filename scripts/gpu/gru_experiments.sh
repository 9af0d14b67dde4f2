#!/bin/bash
# timing experiments on the GRU step kernel (results are WRONG under DP_GRU_DEBUG; timing only)
for d in 0 1 2 3; do
  echo "== DP_GRU_DEBUG=$d"
  DP_GRU_DEBUG=$d timeout 200 python bench.py --steps 3 --warmup 3 --no-cpu-baseline 2>&1 | python -c "
import sys,json
for ln in sys.stdin:
    if ln.startswith('{'):
        d=json.loads(ln); print('value',d['value'],'gru ms/launch',d['roofline']['ms_per_launch'],'TF',d['roofline']['achieved'])
"
done
