#!/bin/bash
mkdir -p gpurun_out
CMD="python bench.py --steps 8 --warmup 3 --no-cpu-baseline --no-x3 --no-parity --no-bam"
for env in "X=1" "DP_GRU_NP=2" "DP_GRU_NP=4" "DP_GRU=wide" "X=2"; do
  env $env timeout 600 $CMD 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$env', round(d['value']/1e6,3), round(d['roofline']['ms_per_launch'],4), round(d['roofline']['frac'],3), d['clocks']['sm_mhz'])"
done
