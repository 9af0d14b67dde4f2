#!/bin/bash
mkdir -p gpurun_out
for v in ctl base w2 w2l9 w4l9 w1l9 w2l9top; do
  timeout 300 python scripts/bgzf_bench.py --kb 6000 --sweep --lib variants/lib_$v.so 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read())['sweep']
print('$v', [(r['blocks'], r['ms']) for r in d if r['blocks'] in (108, 2368, 4736, 5328, 5443)])"
done
