"""TEST INFRASTRUCTURE ONLY -- numpy restatement of the reference model forward and decode.

Restates /root/reference/src/modules/model.py (Model.forward :137-156, residual_layer :93-105, rnn_layer :124-135)
and the decode loop of /root/reference/src/test.py (:59-89) in float64 numpy, LAYER BY LAYER (no operator folding), so
that the folded tensor-core path of the product is checked against the reference's own schedule.  Pinned against the
unmodified reference (oracle/make_model_golden.py -> tests/golden/model_seed7.npz; tests/test_model_oracle.py).
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this.
"""
from __future__ import annotations

from typing import Dict

import numpy as np

BN_EPS = 1e-5


def _conv1x3(x, w, b=None):
    """Cross-correlation along the last axis, zero padding 1 (nn.Conv2d(.., (1,3), padding=(0,1)), model.py:74-77).
    x [R, Cin, W]; w [Cout, Cin, 1, 3] -> [R, Cout, W]"""
    R, Cin, W = x.shape
    xp = np.zeros((R, Cin, W + 2), dtype=x.dtype)
    xp[:, :, 1:W + 1] = x
    out = np.zeros((R, w.shape[0], W), dtype=x.dtype)
    for k in range(3):
        out += np.matmul(w[None, :, :, 0, k], xp[:, :, k:k + W])       # batched [Cout,Cin] @ [Cin,W] (BLAS)
    if b is not None:
        out += b[None, :, None]
    return out


def _conv1x1(x, w):
    return np.matmul(w[None, :, :, 0, 0], x)


def cnn_rows(state: Dict[str, np.ndarray], rows_u8: np.ndarray, dtype=np.float64) -> np.ndarray:
    """The five residual blocks on image rows.  rows_u8 [R, W, C_in] uint8 -> [R, 256, W].
    Kernel height is 1, so each image row goes through the CNN independently of the other two rows of its forward
    call (model.py:72-77); ToTensor's /255 (dataset.py:52) is applied here."""
    s = {k: np.asarray(v, dtype=dtype) for k, v in state.items() if not k.startswith("rnns.")}
    scale = s["batchNorm.weight"] / np.sqrt(s["batchNorm.running_var"] + BN_EPS)
    shift = s["batchNorm.bias"] - s["batchNorm.running_mean"] * scale

    def bn(v):                               # model.py:70, eval mode
        return v * scale[None, :, None] + shift[None, :, None]

    x = np.transpose(rows_u8.astype(dtype) / 255.0, (0, 2, 1))      # [R, C_in, W]
    for layer in range(5):                   # model.py:138-142
        if layer == 0:                       # model.py:94-97 with batch_norm_flag=True
            skip = bn(_conv1x1(x, s["incpConv0.weight"]))
            c1 = bn(_conv1x3(x, s["conv0.weight"]))
        else:
            skip = _conv1x1(x, s["incpConv1.weight"])
            c1 = bn(_conv1x3(x, s["conv1.weight"]))
        c2 = bn(_conv1x3(c1, s["conv2.weight"], s["conv2.bias"]))    # model.py:101
        c3 = _conv1x3(c2, s["conv3.weight"], s["conv3.bias"])        # model.py:102
        x = skip + c3                                                # model.py:104
    return x


def _sigmoid(v):
    return 1.0 / (1.0 + np.exp(-v))


def gru_bidir(x, h0, wih, whh, bih, bhh, wih_r, whh_r, bih_r, bhh_r):
    """torch.nn.GRU(bidirectional=True), one layer, gate order (r, z, n).  x [T, N, H]; h0 [2, N, H].
    Returns (out [T, N, 2H], h_n [2, N, H])."""
    T, N, H = x.shape

    def cell(xt, h, Wi, Wh, bi, bh):
        gi = xt @ Wi.T + bi
        gh = h @ Wh.T + bh
        r = _sigmoid(gi[:, :H] + gh[:, :H])
        z = _sigmoid(gi[:, H:2 * H] + gh[:, H:2 * H])
        n = np.tanh(gi[:, 2 * H:] + r * gh[:, 2 * H:])
        return (1.0 - z) * n + z * h

    out = np.zeros((T, N, 2 * H), dtype=x.dtype)
    h = h0[0]
    for t in range(T):
        h = cell(x[t], h, wih, whh, bih, bhh)
        out[t, :, :H] = h
    hf = h
    h = h0[1]
    for t in range(T - 1, -1, -1):
        h = cell(x[t], h, wih_r, whh_r, bih_r, bhh_r)
        out[t, :, H:] = h
    return out, np.stack([hf, h])


def head(state: Dict[str, np.ndarray], feat: np.ndarray, dtype=np.float64, return_y: bool = False):
    """cnn_fc_rnn + 3 GRUs + fc1 (model.py:151-155).  feat [T, N, C*W] -> logits [N, T, n_class]."""
    g = lambda k: np.asarray(state[k], dtype=dtype)
    y = feat @ g("cnn_fc_rnn.weight").T + g("cnn_fc_rnn.bias")
    x = y
    T, N, H = x.shape
    hidden = np.zeros((2, N, H), dtype=dtype)       # init_hidden (model.py:114-116); only the first 2 slices ever matter
    k = 0
    while "rnns.%d.weight_ih_l0" % k in state:      # model.py:124-135: hidden is chained from layer to layer
        p = "rnns.%d." % k
        out, hidden = gru_bidir(x, hidden, g(p + "weight_ih_l0"), g(p + "weight_hh_l0"), g(p + "bias_ih_l0"),
                                g(p + "bias_hh_l0"), g(p + "weight_ih_l0_reverse"), g(p + "weight_hh_l0_reverse"),
                                g(p + "bias_ih_l0_reverse"), g(p + "bias_hh_l0_reverse"))
        x = out[:, :, :H] + out[:, :, H:]            # sum of directions (model.py:131-133)
        k += 1
    logits = x @ g("fc1.weight").T + g("fc1.bias")
    logits = np.transpose(logits, (1, 0, 2))         # model.py:155
    return (logits, np.transpose(y, (1, 0, 2))) if return_y else logits    # y as [N, T, hidden]


def feat_from_cnn(cnn3: np.ndarray) -> np.ndarray:
    """The reinterpreting view of model.py:144-149.  cnn3 [N, C, 3, W] (contiguous) -> [3, N, C*W]:
    x.view(N, C, W, 3).view(N, C*W, 3).transpose(1,2).transpose(0,1)."""
    N, C, T, W = cnn3.shape
    v = np.ascontiguousarray(cnn3).reshape(N, C, W, T).reshape(N, C * W, T)
    return np.ascontiguousarray(np.transpose(v, (2, 0, 1)))


def forward_image(state: Dict[str, np.ndarray], img_u8: np.ndarray, calls=None, dtype=np.float64, return_y=False):
    """All forward calls of test.py:59-66 for ONE image [H, W, C_in] uint8: call r sees rows r, r+1, r+2.
    Returns logits [n_calls, 3, n_class] (and y [n_calls, 3, hidden])."""
    H = img_u8.shape[0]
    if calls is None:
        calls = list(range(H - 2))
    need = sorted({r + s for r in calls for s in range(3)})
    cnn = cnn_rows(state, img_u8[need], dtype)                        # [len(need), 256, W]
    idx = {r: i for i, r in enumerate(need)}
    cnn3 = np.stack([np.stack([cnn[idx[r + s]] for s in range(3)], axis=1) for r in calls])   # [n, C, 3, W]
    feat = feat_from_cnn(cnn3)
    return head(state, feat, dtype, return_y)


def argmax_first(logits: np.ndarray) -> np.ndarray:
    """torch.topk(1)[1] (test.py:68): index of the first maximum."""
    return np.argmax(logits, axis=-1).astype(np.uint8)


def vote(pred: np.ndarray) -> np.ndarray:
    """test.py:69-89 for one image.  pred [n_calls = H-2, 3] -> calls int8 [H] (-1 where nothing is scored):
    row r >= 2 is voted by (call r-2, t=2), (call r-1, t=1), (call r, t=0); most_common (:23-24) returns the mode, and
    on three distinct votes max(set(l), key=l.count) yields the first element of the set of small ints = the smallest."""
    n = pred.shape[0]
    H = n + 2
    out = np.full(H, -1, dtype=np.int8)
    for row in range(2, n):
        v = [int(pred[row - 2, 2]), int(pred[row - 1, 1]), int(pred[row, 0])]
        out[row] = max(sorted(set(v)), key=v.count)
    return out
