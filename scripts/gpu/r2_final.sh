#!/bin/bash
# final verification of the tree: GPU suite, smoke, both bench arms
mkdir -p gpurun_out
echo "== pytest -m gpu"; timeout 2400 python -m pytest tests -x -q -m gpu 2>&1 | tail -2
echo "== smoke"; timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
echo "== bench"; timeout 1500 python bench.py > gpurun_out/r2_final_bench.json 2> gpurun_out/r2_final_bench.err; echo rc=$?; tail -c 300 gpurun_out/r2_final_bench.err
python -c "
import json; d=json.load(open('gpurun_out/r2_final_bench.json'))
print({k:d.get(k) for k in ('metric','value','unit','ms_per_step','gpu_launches','dtype')}); print('e2e', d['e2e']['value'], 'bam', d.get('e2e_from_bam',{}).get('value'), 'x3', d.get('value_fp16x3'))
print('roofline', d['roofline']['frac'], d['roofline']['ms_per_launch'], d['roofline']['traffic'], d['roofline']['traffic_source'][:40]); print('cpu', d['cpu_baseline']['value'], d['clocks'])"
echo "== reference arm"; timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2_final_reference.json 2>/dev/null; echo rc=$?; cut -c1-300 gpurun_out/r2_final_reference.json
