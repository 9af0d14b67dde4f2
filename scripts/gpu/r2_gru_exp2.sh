#!/bin/bash
mkdir -p gpurun_out; : > gpurun_out/r2_gru_exp2.log
run() { echo "== $*" >> gpurun_out/r2_gru_exp2.log; env "$@" timeout 200 python scripts/gru_micro.py >> gpurun_out/r2_gru_exp2.log 2>&1; }
for epi in 1 2; do
  DP_GRU_EPI=$epi timeout 300 python -m pytest tests/test_model_gpu.py -q -x -k "front_and_logits or batch_composition" 2>&1 | tail -2
done
run DP_GRU_EPI=0
run DP_GRU_EPI=1
run DP_GRU_EPI=2
run DP_GRU_EPI=0 DP_ACT=fast
run DP_GRU_EPI=2 DP_ACT=fast
run DP_GRU_EPI=2 DP_GRU_DEBUG2=2
run DP_GRU_EPI=1 DP_GRU_DEBUG2=2
python - <<'PY'
import json
for ln in open('gpurun_out/r2_gru_exp2.log'):
    if ln.startswith('=='): print(ln.strip(), end='  ')
    elif ln.startswith('{'):
        d=json.loads(ln); print('-> %.4f ms/launch' % d['gru_ms_per_launch'])
    elif 'rror' in ln: print(ln.strip()[:200])
PY
