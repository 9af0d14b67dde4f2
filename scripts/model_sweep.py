"""GPU-box measurement of BASELINE.json configs[3]: model-only sweep of the call batch.

Random legal pileup rows (RGB in {0, 255}, alpha from the quality range, filler to the right; seed 3) go through
`_Engine.forward_rows` (u8->bf16, folded CNN+Linear GEMM, 9 GRU step launches, fc1+argmax) at call batches 64 ... 8192 and at the
bench's 76 800; CUDA events on the launching stream, 3 warm-up + 20 timed repetitions per size, inputs rotated over buffers
larger than L2 where the batch allows.  FLOPs are those of the schedule that runs (7.2 M folded front + 54.0 M
GRU per call).  Prints one JSON object per batch and writes gpurun_out/model_sweep.json.

    python scripts/model_sweep.py [precision]
"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from deepore_b200 import model, modelio                       # noqa: E402

FLOP_PER_CALL = 2.0 * 2400 * 1500 + 9 * 2 * (2.0 * 500 * 1500) * 2 - 2 * (2.0 * 500 * 1500) + 2.0 * 500 * 3 * 3


def peaks():
    try:
        d = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(d.get("bf16_tflops_sustained", d["bf16_tflops"]))
    except Exception:
        return 1400.0


def main():
    prec = sys.argv[1] if len(sys.argv) > 1 else "bf16"
    rng = np.random.default_rng(3)
    m = model.from_state(modelio.seeded_state(7), "cuda", prec)
    sizes = [64, 128, 256, 512, 1024, 2048, 4096, 8192, 76800]
    eng = m.engine(max(sizes), "cuda")
    peak = peaks()
    out = []
    for n in sizes:
        nbuf = max(1, min(8, int(200e6 // (n * 800)) + 1))        # rotate inputs: > 126 MB in flight when n allows
        bufs = []
        for _ in range(nbuf):
            img = np.zeros((n, 200, 4), np.uint8)
            fill = int(rng.integers(20, 120))
            img[:, :fill, :3] = rng.integers(0, 2, size=(n, fill, 3)) * 255
            img[:, :fill, 3] = rng.integers(0, 256, size=(n, fill))
            img[:, fill:, 3] = 255
            bufs.append(torch.from_numpy(img).cuda())
        pred = torch.empty((n, 3), dtype=torch.uint8, device="cuda")
        for i in range(3):
            eng.forward_rows(bufs[i % nbuf], want_logits=False, want_pred=True, pred=pred)
        torch.cuda.synchronize()
        reps = 20
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for i in range(reps):
            eng.forward_rows(bufs[i % nbuf], want_logits=False, want_pred=True, pred=pred)
        b.record()
        torch.cuda.synchronize()
        ms = a.elapsed_time(b) / reps
        tf = n * FLOP_PER_CALL / (ms * 1e-3) / 1e12
        rec = {"call_batch": n, "precision": prec, "ms_per_forward": ms, "calls_per_s": n / (ms * 1e-3),
               "tflops_as_executed": tf, "frac_of_sustained_bf16_peak": tf / peak, "launches_per_forward": 12}
        print(json.dumps(rec), flush=True)
        out.append(rec)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(out, open(os.path.join(ROOT, "gpurun_out", "model_sweep_%s.json" % prec), "w"), indent=1)


if __name__ == "__main__":
    main()
