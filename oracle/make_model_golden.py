"""TEST INFRASTRUCTURE ONLY -- generates tests/golden/model_seed7.npz by running the UNMODIFIED reference model
(/root/reference/src/modules/model.py) on CPU fp32 over pileup images of the bundled NA12878 fixture, following the
loop of /root/reference/src/test.py:51-89 (which itself cannot be imported here: it needs torchnet).

Run in the build container (where /root/reference is mounted):   python oracle/make_model_golden.py
Weights: deepore_b200.modelio.seeded_state(7) (numpy PCG64 -> identical on every box); no checkpoint ships with the
reference.  Parity configuration: Model(input_channel=4, coverage_depth=200), hidden = zeros(2, N, 500) (the
reference's init_hidden returns (6,N,500), which torch >= 0.4 rejects; torch 0.3 read slices 0 and 1 -- SURVEY.md 0.2).
"""
import os
import sys

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
WINDOWS = [0, 57, 131, 259]       # bundled windows (SNP and indel sites)


def most_common(lst):             # test.py:23-24, votes normalised to int (0-dim tensors hash by id on modern torch)
    return max(sorted(set(lst)), key=lst.count)


def main():
    mm = rh.import_reference_model()
    torch.manual_seed(0)
    torch.set_num_threads(os.cpu_count() or 1)
    model = mm.Model(input_channel=4, coverage_depth=200)
    state = modelio.seeded_state(SEED)
    modelio.load_into_module(model, state)
    model.eval()
    ys = []
    model.cnn_fc_rnn.register_forward_hook(lambda m, i, o: ys.append(o.detach().numpy().copy()))

    g = gu.load("bundled_na12878")
    img = np.ascontiguousarray(g.img[WINDOWS])                               # [N,150,200,4]
    images = (torch.from_numpy(img).permute(0, 3, 1, 2).float() / 255.0).contiguous()   # ToTensor (test.py:28)
    N, H = images.size(0), images.size(2)
    seq_len = 3
    logits, preds, calls = [], [], np.full((N, H), -1, dtype=np.int8)
    prediction_stack = []
    with torch.no_grad():
        for row in range(0, H):                                              # test.py:59-89
            if row + seq_len > H:
                continue
            x = images[:, :, row:row + seq_len, :]
            hidden = torch.zeros(2, N, 500)
            out = model(x, hidden)                                           # (N,3,3)
            logits.append(out.numpy().copy())
            p = out.topk(1)[1]                                               # (N,3,1)
            preds.append(p[:, :, 0].numpy().astype(np.uint8))
            prediction_stack.append(p)
            if row + 1 >= seq_len:
                for i in range(N):
                    pr, k = [], seq_len - 1
                    for j in range(len(prediction_stack)):
                        pr.append(int(prediction_stack[j][i][k][0]))
                        k -= 1
                    calls[i, row] = most_common(pr)
                prediction_stack.pop(0)
    logits = np.stack(logits, axis=1)                                        # [N,148,3,3]
    preds = np.stack(preds, axis=1)                                          # [N,148,3]
    y = np.stack(ys, axis=0)                                                 # [148,3,N,500]
    y0 = np.ascontiguousarray(np.transpose(y[:, :, 0, :], (0, 1, 2)))        # window 0: [148,3,500]
    out = os.path.join(ROOT, "tests", "golden", "model_seed%d.npz" % SEED)
    np.savez_compressed(out, seed=SEED, windows=np.asarray(WINDOWS), img=img, logits=logits.astype(np.float32),
                        pred=preds, calls=calls, y_window0=y0.astype(np.float32))
    print("wrote", out, "logits", logits.shape, "class histogram of calls:",
          np.bincount(calls[calls >= 0].astype(np.int64), minlength=3))
    marg = np.sort(logits, axis=-1)
    print("min top-2 margin", float((marg[..., -1] - marg[..., -2]).min()))


if __name__ == "__main__":
    main()
