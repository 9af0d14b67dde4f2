#!/usr/bin/env python
"""How many windows leave the fast encoder kernel for the exact general one (dp_encode_fallback_count)?  Bundled NA12878
windows (260, Illumina 101 bp, 51x) and the synthetic configs[1] / configs[2] workloads.  Windows fall back when they touch
a read outside the fast path's envelope (ops other than M/I/D/S, S inside the CIGAR, I next to I or S, zero-length ops,
a base outside ACGTN), hold more reads than the kernel's slots, or have an illegal reference character.
Note on =/X CIGARs (minimap2 --eqx): in the reference '=' and 'X' advance neither the reference nor the read
(Pileup.py:47-51, deltaRef/deltaRead tables), so such a read piles all its entries on one column -- the exact kernel
reproduces that, a fast path for it would only make a meaningless image faster.
    python scripts/enc_fallback_report.py > gpurun_out/r2_encoder_fallback.json
"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import golden_util as gu                                          # noqa: E402
from deepore_b200.pileup import Encoder, ShardBuffers             # noqa: E402
from deepore_b200.workload import SyntheticWorkload, make_shard   # noqa: E402


def count(shard):
    sb = ShardBuffers(shard, "cuda").upload()
    enc = Encoder(device="cuda")
    _, _, _, status = enc.encode(sb)
    torch.cuda.synchronize()
    return {"windows": shard.n_windows, "fallback_windows": enc.fallback_count(), "status_nonzero": int((status != 0).sum().item()),
            "reads": shard.reads.n_reads}


out = {}
g = gu.load("bundled_na12878")
out["bundled NA12878 chr3:100k-200k, 260 windows at the confident VCF records"] = count(make_shard(g.reads, g.ref, g.ref_base, g.win_qs))
out["configs[1]: synthetic 30x nanopore-length reads, 4096 windows"] = count(SyntheticWorkload().shards(1000, 4096, 1)[0])
out["configs[2]: synthetic 100x, 2048 windows"] = count(SyntheticWorkload(ref_len=1_000_000, coverage=100.0).shards(500, 2048, 1)[0])
print(json.dumps(out, indent=1))
