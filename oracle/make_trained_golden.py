"""TEST INFRASTRUCTURE ONLY -- SURVEY.md 8(f4), the minimum slice: a committed, seeded recipe that takes SGD steps on the 260
bundled NA12878 windows with the UNMODIFIED reference model on CPU, so that a second set of call goldens comes from a
checkpoint that has seen the data (tests/golden/bundled_calls_trained.npz) instead of from seeded random weights only.

The loop is /root/reference/src/train.py:119-181 restated: `nn.CrossEntropyLoss`, `torch.optim.SGD(lr=0.001)`, model in
training mode (BatchNorm on batch statistics, running statistics updated), shuffled batches, one optimiser step per image row
(`x = images[:, :, row:row+3, :]`, `y = labels[:, row:row+3]`), the same two random skip rules (:143-146; with the fixture's
labels -- '1' everywhere but 195 rows -- neither ever fires).  Two things differ, both forced:
  * the optimiser is given only the registered parameters of at most SMALL elements (BatchNorm weight / bias, incpConv0 and
    conv0 weights, conv biases, cnn_fc_rnn.bias, fc1): the full state is 140 MB and cannot be a test fixture, the trained
    small tensors (+ the BatchNorm running statistics) are 40 kB.  Everything else stays at modelio.seeded_state(7).  (The
    reference's own `model.parameters()` never contains the GRUs either: `rnns` is a plain Python list, model.py:124-135.)
  * `hidden` is zeros(2, N, 500) (what rnn_layer consumes; init_hidden's 6 x N x 500 is rejected by today's nn.GRU), and
    oneDNN convolutions are switched off while training (their channels-last outputs break model.py:145's `.view`).

Then the loop of /root/reference/src/test.py:51-89 (eval mode) over all 260 windows, as oracle/make_calls_golden.py does.

    python oracle/make_trained_golden.py [--epochs 1] [--batch 26] [--windows 260]      (build container, ~1 h of CPU)
"""
import argparse
import os
import random
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
SMALL = 4096                      # largest tensor (elements) the optimiser may touch
OUT = os.path.join(ROOT, "tests", "golden", "bundled_calls_trained.npz")


def most_common(lst):             # test.py:23-24
    return max(sorted(set(lst)), key=lst.count)


def train(model, images, labels, epochs, batch_size, log):
    criterion = torch.nn.CrossEntropyLoss()                                          # train.py:120
    trainable = []
    for name, p in model.named_parameters():
        p.requires_grad_(p.numel() <= SMALL)
        if p.requires_grad:
            trainable.append(name)
    optimizer = torch.optim.SGD([p for p in model.parameters() if p.requires_grad], lr=0.001)   # train.py:121
    seq_len, iteration_jump = 3, 1
    gen = torch.Generator().manual_seed(SEED)
    random.seed(SEED)
    steps, t0 = 0, time.time()
    model.train()
    for epoch in range(epochs):
        order = torch.randperm(images.size(0), generator=gen)                        # DataLoader(shuffle=True)
        total_loss, total_images = 0.0, 0
        for b0 in range(0, images.size(0), batch_size):
            idx = order[b0:b0 + batch_size]
            im, lb = images[idx], labels[idx]
            hidden = torch.zeros(2, im.size(0), 500)
            for row in range(0, im.size(2), iteration_jump):                         # train.py:136-168
                if row + seq_len > im.size(2):
                    continue
                x = im[:, :, row:row + seq_len, :]
                y = lb[:, row:row + seq_len]
                total_variation = int(torch.sum(y))
                if total_variation == 0 and random.uniform(0, 1) * 100 > 5:
                    continue
                elif random.uniform(0, 1) < total_variation / batch_size < 0.02:
                    continue
                optimizer.zero_grad()
                outputs = model(x, hidden)
                loss = criterion(outputs.contiguous().view(-1, 3), y.contiguous().view(-1))
                loss.backward()
                optimizer.step()
                total_images += batch_size
                total_loss += float(loss.detach())
                steps += 1
            log("epoch %d windows %d..%d: mean loss %.5f, %d steps, %.0f s"
                % (epoch, b0, b0 + im.size(0) - 1, total_loss / max(1, total_images) * batch_size, steps, time.time() - t0))
    return trainable, steps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--epochs", type=int, default=1)
    ap.add_argument("--batch", type=int, default=26)
    ap.add_argument("--windows", type=int, default=260, help="train on the first N windows (the goldens cover all 260)")
    ap.add_argument("--out", default=OUT)
    args = ap.parse_args()

    def log(msg):
        print(msg, flush=True)

    mm = rh.import_reference_model()
    torch.set_num_threads(os.cpu_count() or 1)
    torch.manual_seed(SEED)
    model = mm.Model(input_channel=4, coverage_depth=200)
    modelio.load_into_module(model, modelio.seeded_state(SEED))
    g = gu.load("bundled_na12878")
    nw, H = g.img.shape[0], g.img.shape[1]
    images = (torch.from_numpy(np.ascontiguousarray(g.img)).permute(0, 3, 1, 2).float() / 255.0).contiguous()   # ToTensor
    labels = torch.tensor([[int(c) for c in str(s)[:H].ljust(H, "0")] for s in g.labels], dtype=torch.long)      # dataset.py:52-54
    before = {k: np.array(v, copy=True) for k, v in modelio.state_from_module(model).items()}   # copies: the views follow the parameters
    torch.backends.mkldnn.enabled = False
    trainable, steps = train(model, images[:args.windows], labels[:args.windows], args.epochs, args.batch, log)
    torch.backends.mkldnn.enabled = True
    model.eval()
    after = modelio.state_from_module(model)
    delta = {}
    for k, v in after.items():
        if not np.array_equal(np.asarray(v), np.asarray(before[k])):
            assert np.asarray(v).size <= SMALL, k
            delta[k] = np.array(v, copy=True)
    log("changed tensors: %s" % sorted(delta))

    seq_len = 3
    pred = np.zeros((nw, H - 2, 3), np.uint8)
    margin = np.zeros((nw, H - 2, 3), np.float32)
    logits = np.zeros((nw, H - 2, 3, 3), np.float32)
    calls = np.full((nw, H), -1, np.int8)
    t0 = time.time()
    with torch.no_grad():
        for b0 in range(0, nw, 52):
            im = images[b0:b0 + 52]
            N = im.size(0)
            for row in range(0, H - seq_len + 1):                                                  # test.py:59-66
                out = model(im[:, :, row:row + seq_len, :].contiguous(), torch.zeros(2, N, 500))
                logits[b0:b0 + N, row] = out.numpy()
                srt = out.sort(dim=-1)[0]
                margin[b0:b0 + N, row] = (srt[..., -1] - srt[..., -2]).numpy()
                pred[b0:b0 + N, row] = out.topk(1)[1][:, :, 0].numpy().astype(np.uint8)
            log("golden windows %d..%d done, %.0f s" % (b0, b0 + N - 1, time.time() - t0))
    for i in range(nw):                                                                            # test.py:68-89
        for row in range(2, H - 2):
            calls[i, row] = most_common([int(pred[i, row - 2, 2]), int(pred[i, row - 1, 1]), int(pred[i, row, 0])])
    lab = labels.numpy()[:, 2:H - 2]
    acc = float((calls[:, 2:H - 2] == lab).mean())
    np.savez_compressed(args.out, seed=SEED, steps=steps, epochs=args.epochs, batch=args.batch, train_windows=args.windows,
                        trainable=np.array(trainable), pred=pred, calls=calls, margin=margin, logits=logits,
                        **{"state/" + k: v for k, v in delta.items()})
    log("wrote %s: %d SGD steps, calls histogram %s, agreement with the labels %.4f, margin min %.3g / 1st percentile %.3g"
        % (args.out, steps, np.bincount(calls[calls >= 0].astype(np.int64), minlength=3), acc, float(margin.min()),
           float(np.percentile(margin, 1))))


if __name__ == "__main__":
    main()
