#!/bin/bash
# GPU-box experiment: GRU step launch time under the kernel's timing switches (DP_GRU_DEBUG / DP_GRU_DEBUG2, tc_gemm.cu).
# Results are wrong by construction in these modes; only the launch time is of interest.
#   DP_GRU_DEBUG : 1 no epilogue, 5 no state loads, 6 no stores, 7 fp32 state store only, 8 no activations, 9 no bias loads
#   DP_GRU_DEBUG2: bit 1 no A loads, bit 2 no MMAs, bit 4 no B loads
for d in "0 0" "7 0" "6 0" "5 0" "8 0" "1 0" "0 2" "0 7" "6 2" "0 0"; do
  set -- $d
  DP_GRU_DEBUG=$1 DP_GRU_DEBUG2=$2 timeout 200 python bench.py --steps 4 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.load(sys.stdin); print('debug $1 mask $2: gru ms/launch', round(d['roofline']['ms_per_launch'],4), 'sm_mhz', d['clocks']['sm_mhz'])" || echo "debug $1 mask $2: failed"
done
