#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 8 --warmup 3 > gpurun_out/r2_final_bench_2gpu.json 2> gpurun_out/b2.err; echo rc=$?
tail -c 300 gpurun_out/b2.err
python -c "
import json; d=json.loads(open('gpurun_out/r2_final_bench_2gpu.json').read().strip().splitlines()[-1])
print({k:d.get(k) for k in ('value','n_gpus','ms_per_step','scaling')}, d['e2e']['value'], d.get('e2e_from_bam',{}).get('value'), d['roofline']['frac'])"
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29534 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 2>/dev/null | tail -1 | cut -c1-200
