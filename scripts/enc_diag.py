"""GPU-box diagnostic: CUDA encoder vs oracle on a synthetic slice; prints the first differing pixels."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import torch
import pileup_oracle as po
from deepore_b200 import synth
from deepore_b200.pileup import Encoder, ShardBuffers
from deepore_b200.workload import make_shard

cov = float(sys.argv[1]) if len(sys.argv) > 1 else 30
ref = synth.random_reference(30_000, seed=5)
rb = synth.simulate_reads(ref, coverage=cov, seed=15)
wq = synth.window_starts(500, 29_500, stride=100)
shard = make_shard(rb, ref, 0, wq)
sb = ShardBuffers(shard, "cuda").upload()
enc = Encoder(device="cuda")
img, rowpos, rowins, status = enc.encode(sb)
torch.cuda.synchronize()
print("fallbacks", enc.fallback_count(), "of", shard.n_windows)
oimg, ost, orp, ori = po.encode_windows(shard.reads, shard.win_qs, shard.win_lo, shard.win_hi, shard.ref,
                                        shard.ref_base, want_rowmap=True)
img = img.cpu().numpy(); rowpos = rowpos.cpu().numpy(); rowins = rowins.cpu().numpy()
print("status equal", np.array_equal(status.cpu().numpy(), ost))
nbad = 0
for i in range(shard.n_windows):
    if np.array_equal(img[i], oimg[i]) and np.array_equal(rowpos[i], orp[i]) and np.array_equal(rowins[i], ori[i]):
        continue
    nbad += 1
    if nbad > 4:
        continue
    d = np.argwhere((img[i] != oimg[i]).any(axis=2))
    print("window", i, "qs", int(shard.win_qs[i]), "differing pixels", len(d), "rowpos eq", np.array_equal(rowpos[i], orp[i]),
          "rowins eq", np.array_equal(rowins[i], ori[i]))
    for (y, x) in d[:12]:
        print("   row", y, "x", x, "col", int(orp[i][y]) - int(shard.win_qs[i]), "ins", int(ori[i][y]), "got", img[i][y, x], "want", oimg[i][y, x])
print("windows differing:", nbad, "of", shard.n_windows)
