#!/bin/bash
mkdir -p gpurun_out
timeout 900 python scripts/model_sweep.py fp16 > gpurun_out/sweep.log 2>&1; tail -3 gpurun_out/sweep.log | cut -c1-400
cp gpurun_out/model_sweep.json gpurun_out/r2_config4_model_sweep_fp16.json 2>/dev/null
timeout 900 python scripts/enc_cov_time.py > gpurun_out/r2_config3_encoder_coverage.json 2>gpurun_out/cov.err; tail -c 600 gpurun_out/r2_config3_encoder_coverage.json
