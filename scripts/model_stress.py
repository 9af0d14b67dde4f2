"""GPU-box stress run (not part of the test suite): CUDA model forward vs the layer-by-layer numpy oracle for several weight
seeds and random legal pileup tensors, the default, the parity-grade and the round-1 precision mode.

    python scripts/model_stress.py [n_seeds] [first_seed]
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import model_oracle as mo                                     # noqa: E402
from deepore_b200 import model, modelio                       # noqa: E402

TOL = {"fp16": 6e-3, "fp16x3": 2e-4, "bf16": 3e-2}      # absolute, logit scales 3 - 5 (1.2e-3 x scale, 1e-4)


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 6
    s0 = int(sys.argv[2]) if len(sys.argv) > 2 else 100
    bad = 0
    for seed in range(s0, s0 + n):
        rng = np.random.default_rng(seed)
        st = modelio.seeded_state(seed)
        nw = int(rng.integers(1, 4))
        img = np.zeros((nw, 150, 200, 4), np.uint8)
        img[..., :3] = rng.integers(0, 2, size=(nw, 150, 200, 3)) * 255
        img[..., 3] = rng.integers(0, 256, size=(nw, 150, 200))
        fill = rng.integers(20, 200)                            # realistic shape: filler to the right of `fill` entries
        img[:, :, fill:, :3] = 0
        img[:, :, fill:, 3] = 255
        calls = sorted(set([0, 1, 146, 147] + [int(x) for x in rng.integers(2, 146, size=4)]))
        want = np.stack([mo.forward_image(st, img[i], calls=calls) for i in range(nw)])
        line = "seed %d, %d windows, %d calls:" % (seed, nw, len(calls))
        for prec in ("fp16x3", "fp16", "bf16"):
            m = model.from_state(st, "cuda", prec)
            logits, _ = m.forward_windows(torch.from_numpy(img).cuda())
            torch.cuda.synchronize()
            err = float(np.abs(logits.cpu().numpy()[:, calls] - want).max())
            ok = err < TOL[prec]
            bad += 0 if ok else 1
            line += "  %s max abs err %.3g %s" % (prec, err, "ok" if ok else "OVER")
            del m
        print(line, "(logit scale %.2f)" % float(np.abs(want).max()), flush=True)
    print("%d seeds, %d over tolerance" % (n, bad))
    sys.exit(1 if bad else 0)


if __name__ == "__main__":
    main()
