"""GPU-box measurement of BASELINE.json configs[2]: the pileup encoder alone on synthetic shards of 30x and
100x coverage (4096 windows per call, window centres every 100 bp; same read generator as configs[1]).  CUDA events on the
launching stream, 3 warm-up + 20 timed calls rotating over shards whose output exceeds L2 (492 MB of pileup tensors per call).
Reports windows/s and algorithmic GB/s (workload.algorithmic_bytes) against the measured HBM peak; writes
gpurun_out/enc_coverage.json.
"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from deepore_b200 import synth                                            # noqa: E402
from deepore_b200.pileup import Encoder, ShardBuffers                     # noqa: E402
from deepore_b200.workload import make_shard, algorithmic_bytes           # noqa: E402

try:
    PEAK = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
except Exception:
    PEAK = 6650.0
NW = 4096
out = []
COVS = [int(x) for x in sys.argv[1:]] or [30, 100]
for cov in COVS:
    ref = synth.random_reference(NW * 100 * 2 + 40_000, seed=1)
    rb = synth.simulate_reads(ref, coverage=cov, seed=2)
    shards = []
    for k in range(2):
        wq = synth.window_starts(10_000 + k * NW * 100, 10_000 + (k + 1) * NW * 100 + 201, stride=100)[:NW]
        shards.append(make_shard(rb, ref, 0, wq))
    sbs = [ShardBuffers(sh, "cuda").upload() for sh in shards]
    enc = Encoder(device="cuda")
    for i in range(3):
        enc.encode(sbs[i & 1])
    torch.cuda.synchronize()
    reps = 20
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(reps):
        enc.encode(sbs[i & 1])
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / reps
    alg = float(np.mean([algorithmic_bytes(sh) for sh in shards]))
    nw = int(np.mean([sh.n_windows for sh in shards]))
    gbs = alg / (ms * 1e-3) / 1e9
    rec = {"coverage": cov, "windows_per_call": nw, "ms_per_call": ms, "windows_per_s": nw / (ms * 1e-3),
           "algorithmic_bytes_per_window": alg / nw, "achieved_gbs": gbs, "hbm_peak_gbs": PEAK, "frac": gbs / PEAK,
           "fallback_windows": int(enc.fallback_count())}
    print(json.dumps(rec), flush=True)
    out.append(rec)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "enc_coverage.json"), "w"), indent=1)
