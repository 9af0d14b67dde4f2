#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29517 scripts/config5_stream.py --chunks 10 --distinct 2 > gpurun_out/r2_config5_v5_8gpu.json 2> gpurun_out/c5_8.err; echo rc=$?
tail -c 400 gpurun_out/c5_8.err; python -c "
import json; d=json.loads(open('gpurun_out/r2_config5_v5_8gpu.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ('n_gpus','positions_per_s','seconds','ingest_ms_per_chunk_side_stream','compute_ms_per_chunk','host_prep_s_untimed')})"
