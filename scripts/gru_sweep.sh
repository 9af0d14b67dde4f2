#!/bin/bash
# GPU-box experiment: GRU step launch time under the kernel's timing switches (DP_GRU_DEBUG / DP_GRU_DEBUG2, tc_gemm.cu).
# Results are wrong by construction in these modes; only the launch time is of interest.
for d in "0 0" "1 0" "1 1" "1 4" "1 5" "1 2" "1 7"; do
  set -- $d
  DP_GRU_DEBUG=$1 DP_GRU_DEBUG2=$2 timeout 200 python bench.py --steps 4 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.load(sys.stdin); print('debug $1 mask $2: gru ms/launch', round(d['roofline']['ms_per_launch'],4))"
done
