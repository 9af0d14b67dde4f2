#!/bin/bash
mkdir -p gpurun_out; : > gpurun_out/r2_gru_exp3.log
run() { echo "== $*" >> gpurun_out/r2_gru_exp3.log; env "$@" timeout 200 python scripts/gru_micro.py >> gpurun_out/r2_gru_exp3.log 2>&1; }
run DP_GRU_STREAM=0
run DP_GRU_STREAM=1
run DP_GRU_STREAM=0
run DP_GRU_STREAM=1
python scripts/bgzf_bench.py --kb 6000 2>/dev/null | python -c "
import json,sys; d=json.load(sys.stdin); print({k:round(v,3) for k,v in d.items() if k.endswith('_ms')})"
python scripts/enc_fallback_report.py > gpurun_out/r2_encoder_fallback.json 2>/dev/null; cat gpurun_out/r2_encoder_fallback.json | tr -d '\n' | cut -c1-700; echo
python - <<'PY'
import json
for ln in open('gpurun_out/r2_gru_exp3.log'):
    if ln.startswith('=='): print(ln.strip(), end='  ')
    elif ln.startswith('{'):
        d=json.loads(ln); print('-> %.4f ms/launch' % d['gru_ms_per_launch'])
    elif 'rror' in ln: print(ln.strip()[:200])
PY
