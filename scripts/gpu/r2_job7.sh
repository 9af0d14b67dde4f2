#!/bin/bash
# bench lines of round 2 + ncu evidence (each ncu pass only after the same command exited 0 without ncu)
mkdir -p gpurun_out
T=${1:-v3}
timeout 300 python __graft_entry__.py --smoke 2>&1 | tail -2
timeout 900 python bench.py > gpurun_out/r2_e2e_${T}_bench.json 2> gpurun_out/r2_e2e_${T}.err; echo "bench rc=$?"; tail -2 gpurun_out/r2_e2e_${T}.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2_reference_arm_${T}_bench.json 2>> gpurun_out/r2_e2e_${T}.err; echo "bench ref rc=$?"
timeout 600 python bench.py --stage encode --steps 40 --warmup 3 --windows 4096 > gpurun_out/r2_encode_${T}_bench.json 2>> gpurun_out/r2_e2e_${T}.err; echo "bench enc rc=$?"
timeout 900 python scripts/bgzf_bench.py --kb 6000 > gpurun_out/r2_bgzf_${T}_bench_6000kb.json 2>> gpurun_out/r2_e2e_${T}.err; echo "bgzf rc=$?"
CMD="python bench.py --steps 2 --warmup 3 --windows 1024 --no-cpu-baseline --no-x3 --no-parity --no-bam"
timeout 300 $CMD > gpurun_out/plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_e2e_${T}_launches.csv $CMD > gpurun_out/ncu_launch.log 2>&1
echo "ncu launch list rc=$?"
timeout 300 $CMD > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gru_step2 -s 20 -c 2 -f -o gpurun_out/r2_gru_full_$T $CMD > gpurun_out/ncu_gru.log 2>&1
echo "ncu gru full rc=$?"
for NW in 2048 4096; do
CMDE="python bench.py --stage encode --steps 2 --warmup 3 --windows $NW --no-cpu-baseline"
timeout 300 $CMDE > /dev/null 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"prep_reads|base_codes|encode_windows_v3" -s 12 -c 3 -f -o gpurun_out/r2_enc_full_${NW}_$T $CMDE > gpurun_out/ncu_enc_full.log 2>&1
echo "ncu enc $NW full rc=$?"
done
CMDB="python scripts/bgzf_bench.py --kb 6000 --reps 1"
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum --clock-control none -k regex:"bgzf_inflate|bam_" -c 40 --csv --log-file gpurun_out/r2_bgzf_${T}_ncu_launches.csv $CMDB > gpurun_out/ncu_bgzf.log 2>&1
echo "ncu bgzf rc=$?"
ls -la gpurun_out | tail -12
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/r2_e2e_v3_bench.json') if l.startswith('{')][0])
print('value %.4g e2e %.4g frac %.3f ms/launch %.4f' % (d['value'], d['e2e']['value'], d['roofline']['frac'], d['roofline']['ms_per_launch']))
for k in ('value_fp16x3','e2e_fp16x3','parity','e2e_from_bam','cpu_baseline','roofline_encoder','output_checked','status_checked'):
    print(k, json.dumps(d.get(k))[:600])
PY
cat gpurun_out/r2_reference_arm_v3_bench.json | cut -c1-300
cat gpurun_out/r2_bgzf_v3_bench_6000kb.json | cut -c1-900
