#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -s -x tests/test_pipeline.py > gpurun_out/r2_job3_pipeline.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_job3_pipeline.log
tail -15 gpurun_out/r2_job3_pipeline.log
timeout 900 python scripts/model_error_attribution.py --only fp16 --out gpurun_out/r2_model_error_attribution_fp16.json > gpurun_out/r2_job3_attr.log 2>&1
cat gpurun_out/r2_job3_attr.log | tail -12
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/r2_bench_v2.json 2> gpurun_out/r2_bench_v2.err
tail -3 gpurun_out/r2_bench_v2.err
python - <<'PY'
import json
for ln in open('gpurun_out/r2_bench_v2.json'):
    if ln.startswith('{'):
        d=json.loads(ln); print('value %.4g e2e %.4g' % (d['value'], d['e2e']['value'])); print(json.dumps(d.get('e2e_from_bam'), indent=1))
PY
