"""Helpers shared by the tests: load a tests/golden/*.npz fixture as (ReadBatch, ref, ref_base, wins, img...)."""
import os
from types import SimpleNamespace

import numpy as np

from deepore_b200.reads import ReadBatch

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
FIXTURES = ["quirks", "deep", "fuzz", "nanopore_small", "bundled_na12878"]


def load(name):
    z = np.load(os.path.join(GOLD, name + ".npz"))
    rb = ReadBatch(contig="chr3", **{k: z[k] for k in
                                    ("pos", "flag", "mapq", "l_seq", "cigar_off", "cigar", "seq_off", "seq4",
                                     "qual_off", "qual")})
    extra = {k: z[k] for k in z.files}
    return SimpleNamespace(reads=rb, ref=z["ref"], ref_base=int(z["ref_base"]), win_qs=z["win_qs"],
                           img=z["img"], status=z["status"], labels=z["labels"], z=extra)
