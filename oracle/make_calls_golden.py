"""TEST INFRASTRUCTURE ONLY -- generates tests/golden/bundled_calls_seed7.npz: the call set of the UNMODIFIED reference
model (/root/reference/src/modules/model.py, torch CPU fp32) over ALL 260 pileup images of the bundled NA12878 fixture,
following the loop of /root/reference/src/test.py:51-89 (forward per image row, top-1, 3-way vote).

Run in the build container (where /root/reference is mounted):   python oracle/make_calls_golden.py
Stored per window: the raw logits (float16-rounded copies would hide the differences the tests count, so float32), the
top-1 class of every (call, t), the vote of every scored row, and the top-2 logit margin of every (call, t) -- the GPU
tests (tests/test_model_gpu.py) count EVERY top-1 / vote that differs and demand that each difference sits where the
reference's own margin is below twice the measured logit error.  Weights: deepore_b200.modelio.seeded_state(7); no checkpoint ships with the reference.
"""
import os
import sys
import time

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import golden_util as gu          # noqa: E402
import ref_harness as rh          # noqa: E402
from deepore_b200 import modelio  # noqa: E402

SEED = 7
BATCH = 52                        # windows per forward call (test.py uses a DataLoader batch the same way)


def most_common(lst):             # test.py:23-24
    return max(sorted(set(lst)), key=lst.count)


def main():
    mm = rh.import_reference_model()
    torch.set_num_threads(os.cpu_count() or 1)
    model = mm.Model(input_channel=4, coverage_depth=200)
    modelio.load_into_module(model, modelio.seeded_state(SEED))
    model.eval()
    g = gu.load("bundled_na12878")
    nw, H = g.img.shape[0], g.img.shape[1]
    seq_len = 3
    pred = np.zeros((nw, H - 2, 3), np.uint8)
    margin = np.zeros((nw, H - 2, 3), np.float32)
    logits = np.zeros((nw, H - 2, 3, 3), np.float32)
    calls = np.full((nw, H), -1, np.int8)
    t0 = time.time()
    with torch.no_grad():
        for b0 in range(0, nw, BATCH):
            img = np.ascontiguousarray(g.img[b0:b0 + BATCH])
            images = (torch.from_numpy(img).permute(0, 3, 1, 2).float() / 255.0).contiguous()      # ToTensor (test.py:28)
            N = images.size(0)
            for row in range(0, H - seq_len + 1):                                                  # test.py:59-66
                out = model(images[:, :, row:row + seq_len, :], torch.zeros(2, N, 500))            # (N,3,3)
                logits[b0:b0 + N, row] = out.numpy()
                srt = out.sort(dim=-1)[0]
                margin[b0:b0 + N, row] = (srt[..., -1] - srt[..., -2]).numpy()
                pred[b0:b0 + N, row] = out.topk(1)[1][:, :, 0].numpy().astype(np.uint8)
            print("windows %d..%d done, %.0f s" % (b0, b0 + N - 1, time.time() - t0), flush=True)
    for i in range(nw):                                                                            # test.py:68-89
        for row in range(2, H - 2):
            calls[i, row] = most_common([int(pred[i, row - 2, 2]), int(pred[i, row - 1, 1]), int(pred[i, row, 0])])
    out = os.path.join(ROOT, "tests", "golden", "bundled_calls_seed%d.npz" % SEED)
    np.savez_compressed(out, seed=SEED, pred=pred, calls=calls, margin=margin, logits=logits)
    print("wrote", out, "class histogram of calls:", np.bincount(calls[calls >= 0].astype(np.int64), minlength=3),
          "min margin", float(margin.min()))


if __name__ == "__main__":
    main()
