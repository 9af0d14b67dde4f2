#!/bin/bash
mkdir -p gpurun_out
CMDB="python scripts/bgzf_bench.py --kb 6000"
timeout 600 $CMDB > gpurun_out/r2_bgzf_v9_bench_6000kb.json 2>/dev/null
python -c "
import json; d=json.load(open('gpurun_out/r2_bgzf_v9_bench_6000kb.json')); print({k:round(v,3) for k,v in d.items() if isinstance(v,float)})"
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum --clock-control none -k regex:"bgzf_inflate|bam_" -c 8 --csv --log-file gpurun_out/r2_bgzf_v9_ncu_launches.csv $CMDB --reps 1 > /dev/null 2>&1
grep bgzf_inflate gpurun_out/r2_bgzf_v9_ncu_launches.csv | head -4 | awk -F'","' '{print $13, $15}'
timeout 900 ncu --set full --import-source on --clock-control none -k regex:bgzf_inflate -c 1 -o gpurun_out/r2_inflate_v9b_full -f $CMDB --reps 1 > /dev/null 2>&1
timeout 600 python scripts/bgzf_bench.py --kb 12000 --sweep 2>/dev/null > gpurun_out/r2_inflate_v9_sweep.json; cat gpurun_out/r2_inflate_v9_sweep.json | cut -c1-900
