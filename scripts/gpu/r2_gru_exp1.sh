#!/bin/bash
# GRU step pipeline experiments: which part of the wide-tile kernels is slow?  (timing switches; results wrong by construction)
mkdir -p gpurun_out; : > gpurun_out/r2_gru_exp1.log
run() { echo "== $*" >> gpurun_out/r2_gru_exp1.log; env "$@" timeout 200 python scripts/gru_micro.py >> gpurun_out/r2_gru_exp1.log 2>&1; }
run DP_GRU=pair
run DP_GRU=pair DP_GRU_DEBUG=1
run DP_GRU=pair DP_GRU_DEBUG2=2
run DP_GRU=pair DP_GRU_DEBUG=1 DP_GRU_DEBUG2=2
run DP_GRU=wide
run DP_GRU=wide DP_GRU_DEBUG=1
run DP_GRU=wide DP_GRU_DEBUG2=2
run DP_GRU=wide DP_GRU_DEBUG=1 DP_GRU_DEBUG2=2
run DP_GRU=stag DP_GRUS_CFG=2
run DP_GRU=stag DP_GRUS_CFG=2 DP_GRU_DEBUG=1
run DP_GRU=stag DP_GRUS_CFG=2 DP_GRU_DEBUG2=2
run DP_GRU=stag DP_GRUS_CFG=2 DP_GRU_DEBUG=1 DP_GRU_DEBUG2=2
python - <<'PY'
import json
for ln in open('gpurun_out/r2_gru_exp1.log'):
    if ln.startswith('=='): print(ln.strip(), end='  ')
    elif ln.startswith('{'):
        d=json.loads(ln); print('-> %.4f ms/launch' % d['gru_ms_per_launch'])
    elif 'Error' in ln or 'error' in ln: print(ln.strip()[:200])
PY
