"""GPU-box experiment: encoder time per window on synthetic shards of a given coverage (CUDA events, 20 repetitions)."""
import sys, os
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from deepore_b200 import synth
from deepore_b200.pileup import Encoder, ShardBuffers
from deepore_b200.workload import make_shard
for cov in (30, 100):
    ref = synth.random_reference(230_000, seed=1)
    rb = synth.simulate_reads(ref, coverage=cov, seed=2)
    wq = synth.window_starts(10_000, 10_000 + 2048 * 100, stride=100)
    sb = ShardBuffers(make_shard(rb, ref, 0, wq), "cuda").upload()
    enc = Encoder(device="cuda")
    for _ in range(3):
        enc.encode(sb)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(20):
        enc.encode(sb)
    b.record(); torch.cuda.synchronize()
    ms = a.elapsed_time(b) / 20
    print("coverage %d: %d windows, %.3f ms per call, %.2f M windows/s, fallback %d" % (cov, len(wq), ms, len(wq) / ms / 1e3, enc.fallback_count()))
