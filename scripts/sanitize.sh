#!/bin/bash
# compute-sanitizer over the byte-work kernels (SURVEY 5 aux hooks): the encoder leans on shared-memory atomics, carved
# pools and cp.async.bulk lifetime rules, the ingest kernels on bounds checks of untrusted input.  memcheck finds
# out-of-bounds / misaligned accesses, racecheck shared-memory hazards between the barriers.  Run on a GPU box:
#     bash scripts/sanitize.sh            (logs: gpurun_out/sanitize_*.log; summaries are copied to profiles/)
# The tensor-core kernels are left out: racecheck does not model mbarrier / TMA / tcgen05 ordering and reports every
# asynchronous-proxy write as a hazard.
set -u
mkdir -p gpurun_out
SAN=${SAN:-/usr/local/cuda/bin/compute-sanitizer}
SEL_ENC='bit_exact_vs_reference_golden or small_image_parameters or empty_inputs or rowmap_matches_oracle'
SEL_BGZF='inflate_bit_exact_vs_zlib or inflate_ragged or corrupt or records_match'
run() {   # tool, name, test file, -k selection
  echo "== $1 $2" | tee -a gpurun_out/sanitize_summary.txt
  timeout 1500 $SAN --tool $1 --error-exitcode 9 --launch-timeout 0 --print-limit 20 \
      python -m pytest "$3" -q -x -k "$4" > gpurun_out/sanitize_$1_$2.log 2>&1
  echo "exit code $?" | tee -a gpurun_out/sanitize_summary.txt
  grep -E "ERROR SUMMARY|RACECHECK SUMMARY|passed|failed" gpurun_out/sanitize_$1_$2.log | tail -4 | tee -a gpurun_out/sanitize_summary.txt
}
: > gpurun_out/sanitize_summary.txt
run memcheck encoder tests/test_encoder_gpu.py "$SEL_ENC"
run memcheck bgzf tests/test_bgzf_gpu.py "$SEL_BGZF"
run racecheck encoder tests/test_encoder_gpu.py "bit_exact_vs_reference_golden and (quirks or fuzz)"
run racecheck bgzf tests/test_bgzf_gpu.py "inflate_bit_exact_vs_zlib"
