"""GPU-box stress run (not part of the test suite): CUDA encoder vs the C oracle on randomised synthetic workloads -- read
length 60 bp .. 30 kb, coverage 3 .. 150, indel rates up to 12 %, heavy soft clipping, N bases in reads and reference,
window strides 1 .. 250.  Prints one line per case and exits non-zero on the first difference.

    python scripts/enc_stress.py [n_cases] [first_seed]
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import pileup_oracle as po                                   # noqa: E402
from deepore_b200 import synth                                # noqa: E402
from deepore_b200.pileup import Encoder, ShardBuffers         # noqa: E402
from deepore_b200.workload import make_shard                  # noqa: E402


def main():
    n_cases = int(sys.argv[1]) if len(sys.argv) > 1 else 40
    seed0 = int(sys.argv[2]) if len(sys.argv) > 2 else 1000
    enc = Encoder(device="cuda")
    bad = 0
    for case in range(n_cases):
        rng = np.random.default_rng(seed0 + case)
        G = int(rng.integers(8_000, 60_000))
        mean_len = float(np.exp(rng.uniform(np.log(60), np.log(30_000))))
        cov = float(np.exp(rng.uniform(np.log(3), np.log(150))))
        if mean_len < 300:
            cov = min(cov, 60.0)                              # short reads: many more reads per window at equal coverage
        kw = dict(coverage=cov, seed=int(rng.integers(1, 1 << 30)), mean_len=mean_len, sigma=float(rng.uniform(0.05, 0.8)),
                  min_len=int(max(30, mean_len / 8)), sub=float(rng.uniform(0, 0.1)), ins=float(rng.uniform(0, 0.12)),
                  dele=float(rng.uniform(0, 0.12)), indel_p=float(rng.uniform(0.3, 0.9)),
                  clip_frac=float(rng.uniform(0, 0.6)), clip_max=int(rng.integers(1, 400)),
                  lowmapq_frac=float(rng.uniform(0, 0.3)), read_n_frac=float(rng.choice([0.0, 0.01])))
        ref = synth.random_reference(G, seed=int(rng.integers(1, 1 << 30)), n_frac=float(rng.choice([0.0, 0.002])))
        rb = synth.simulate_reads(ref, **kw)
        stride = int(rng.choice([1, 7, 50, 100, 250]))
        wq = synth.window_starts(300, G - 300, stride=stride)
        if wq.shape[0] > 600:
            wq = wq[:600]
        shard = make_shard(rb, ref, 0, wq)
        sb = ShardBuffers(shard, "cuda").upload()
        img, rowpos, rowins, status = enc.encode(sb)
        torch.cuda.synchronize()
        nfb = enc.fallback_count()
        oimg, ost, orp, ori = po.encode_windows(shard.reads, shard.win_qs, shard.win_lo, shard.win_hi, shard.ref,
                                                shard.ref_base, want_rowmap=True)
        ok = (np.array_equal(status.cpu().numpy(), ost) and np.array_equal(img.cpu().numpy(), oimg)
              and np.array_equal(rowpos.cpu().numpy(), orp) and np.array_equal(rowins.cpu().numpy(), ori))
        depth = float(np.mean(shard.win_hi - shard.win_lo)) if shard.n_windows else 0.0
        print("case %3d seed %d: G %d len %.0f cov %.1f stride %d windows %d reads %d depth %.0f fallback %d status!=0 %d -> %s"
              % (case, seed0 + case, G, mean_len, cov, stride, shard.n_windows, rb.n_reads, depth, nfb,
                 int((ost != 0).sum()), "ok" if ok else "DIFFERENT"), flush=True)
        if not ok:
            bad += 1
            diff = [i for i in range(shard.n_windows) if not np.array_equal(img[i].cpu().numpy(), oimg[i])]
            print("   windows that differ:", diff[:20], flush=True)
    print("%d cases, %d different" % (n_cases, bad))
    sys.exit(1 if bad else 0)


if __name__ == "__main__":
    main()
