#!/bin/bash
mkdir -p gpurun_out
timeout 900 python scripts/bgzf_bench.py --kb 6000 --sweep 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read())['sweep']
print([(r['blocks'], r['ms']) for r in d])"
timeout 600 python -m pytest tests -m gpu -q -k "bgzf" 2>&1 | tail -1
