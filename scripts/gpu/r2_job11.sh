#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -k "bgzf or pipeline or validation_pass" 2>&1 | tail -2
CMDB="python scripts/bgzf_bench.py --kb 6000"
timeout 600 $CMDB > gpurun_out/r2_bgzf_v8_bench_6000kb.json 2>/dev/null
python -c "
import json; d=json.load(open('gpurun_out/r2_bgzf_v8_bench_6000kb.json')); print({k:round(v,3) for k,v in d.items() if isinstance(v,float)})"
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum --clock-control none -k regex:"bgzf_inflate|bam_" -c 8 --csv --log-file gpurun_out/r2_bgzf_v8_ncu_launches.csv $CMDB --reps 1 > /dev/null 2>&1
grep bgzf_inflate gpurun_out/r2_bgzf_v8_ncu_launches.csv | head -4 | awk -F'","' '{print $13, $15}'
