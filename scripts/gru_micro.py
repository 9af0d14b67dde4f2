#!/usr/bin/env python
"""GRU step kernel micro-benchmark: ms per launch of the model's 9 GRU step launches at the bench's launch size (76 800 rows),
CUDA events on the launching stream (dp_model_profile), for the kernel variant / timing switches in the environment
(DP_GRU, DP_GRUS_CFG, DP_GRU_NP, DP_GRU_DEBUG, DP_GRU_DEBUG2 -- results are WRONG under the debug switches, only the time
matters).  One process per configuration: the switches are read once.

    DP_GRU=stag python scripts/gru_micro.py [precision] [windows]
"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from deepore_b200 import model, modelio                       # noqa: E402


def main():
    prec = sys.argv[1] if len(sys.argv) > 1 else "fp16"
    nw = int(sys.argv[2]) if len(sys.argv) > 2 else 512
    rng = np.random.default_rng(3)
    m = model.from_state(modelio.seeded_state(7), "cuda", prec)
    H = 150
    eng = m.engine(nw * H, "cuda")
    bufs = []
    for _ in range(3):                                        # rotate inputs: 3 x 61 MB of image rows
        img = np.zeros((nw * H, 200, 4), np.uint8)
        img[:, :40, :3] = rng.integers(0, 2, size=(nw * H, 40, 3)) * 255
        img[:, :40, 3] = rng.integers(0, 256, size=(nw * H, 40))
        img[:, 40:, 3] = 255
        bufs.append(torch.from_numpy(img).cuda())
    pred = torch.empty((nw * H, 3), dtype=torch.uint8, device="cuda")
    for i in range(3):
        eng.forward_rows(bufs[i % 3], want_logits=False, pred=pred)
    torch.cuda.synchronize()
    eng.profile(True)
    reps = 12
    for i in range(reps):
        eng.forward_rows(bufs[i % 3], want_logits=False, pred=pred)
    prof = eng.profile_read()
    g_ms, g_n = prof["gru_step"]
    f_ms, f_n = prof["fold_gemm"]
    env = {k: v for k, v in os.environ.items() if k.startswith("DP_GRU")}
    flop = nw * (H - 2) * (8 * 2 * (2.0 * 500 * 1500) * 2 + 2 * (2.0 * 500 * 1500)) / 9
    print(json.dumps({"env": env, "precision": prec, "rows": nw * H, "gru_ms_per_launch": g_ms / g_n, "launches": g_n,
                      "gru_tflops_algorithmic": flop / (g_ms / g_n * 1e-3) / 1e12, "fold_ms": f_ms / f_n}))


if __name__ == "__main__":
    main()
