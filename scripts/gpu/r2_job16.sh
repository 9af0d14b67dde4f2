#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q -k "bgzf or pipeline" 2>&1 | tail -1
timeout 900 python scripts/bgzf_bench.py --kb 6000 --sweep 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read())['sweep']
print([(r['blocks'], r['ms']) for r in d])"
CMDB="python scripts/bgzf_bench.py --kb 6000"
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum --clock-control none -k regex:"bgzf_inflate" -c 2 --csv --log-file gpurun_out/r2_bgzf_v9_ncu_launches.csv $CMDB --reps 1 > /dev/null 2>&1
grep bgzf_inflate gpurun_out/r2_bgzf_v9_ncu_launches.csv | head -4 | awk -F'","' '{print $13, $15}'
