#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -k "windows or pipeline or encoder_gpu and quirks" 2>&1 | tail -3
timeout 600 python bench.py --no-cpu-baseline --no-x3 > gpurun_out/r2_e2e_v4_bench.json 2> gpurun_out/r2_e2e_v4.err; echo "rc=$?"; tail -3 gpurun_out/r2_e2e_v4.err
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/r2_e2e_v4_bench.json') if l.startswith('{')][0])
b=d['e2e_from_bam']; print('value %.4g e2e %.4g bam %.4g ingest %.1f ms compute %.1f ms' % (d['value'], d['e2e']['value'], b['value'], b['ingest_ms_per_chunk'], b['compute_ms_per_chunk']))
PY
