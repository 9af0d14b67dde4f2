#!/usr/bin/env python
"""Where does the logits error of the tensor-core model path come from?  ("next round" item 1a.)

Runs the CUDA forward over the golden windows with ONE error source at a time switched to its accurate variant and
prints, per configuration: logits max-abs / rms against the UNMODIFIED reference model (torch CPU fp32 logits stored in
tests/golden/bundled_calls_seed7.npz for all 260 bundled windows, tests/golden/model_seed7.npz for 4 of them), the number
of top-1 classes and 3-way votes that differ over ALL 260 x 148 x 3 / 37 960 of them, the largest reference margin at
which a top-1 differs, and the time of one forward over the 260 windows.

Switches (read by dp_model_create, csrc/model.cu): precision code (bf16 / fp16 operands, one pass or split hi+lo three
passes), DP_FOLD_SPLIT (folded CNN+Linear weights hi+lo), DP_GRU_SPLIT (bit mask of GRU layers that run the split
passes), DP_ACT (fast = MUFU.TANH, accurate = exp-based), DP_GRU_STATE (f32 | bf16: recurrent state kept in fp32, or
only in the 16-bit operand rows).

    python scripts/model_error_attribution.py [--out gpurun_out/model_error_attribution.json]
"""
import argparse
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
GOLD = os.path.join(ROOT, "tests", "golden")

SWITCHES = ("DP_FOLD_SPLIT", "DP_GRU_SPLIT", "DP_ACT", "DP_GRU_STATE")

CONFIGS = [
    # name, precision, env
    ("bf16 (round-1 default: bf16 operands, MUFU.TANH, fp32 state)", "bf16", {}),
    ("bf16 + folded weights hi+lo", "bf16", {"DP_FOLD_SPLIT": "1"}),
    ("bf16 + exp-based activations", "bf16", {"DP_ACT": "accurate"}),
    ("bf16 + GRU layer 0 split", "bf16", {"DP_GRU_SPLIT": "1"}),
    ("bf16 + GRU layer 1 split", "bf16", {"DP_GRU_SPLIT": "2"}),
    ("bf16 + GRU layer 2 split", "bf16", {"DP_GRU_SPLIT": "4"}),
    ("bf16 + all GRU layers split", "bf16", {"DP_GRU_SPLIT": "7"}),
    ("bf16 + 16-bit state only", "bf16", {"DP_GRU_STATE": "bf16"}),
    ("bf16x3 (all split, exp-based)", "bf16x3", {}),
    ("fp16 with MUFU.TANH (fp16 operands, fp32 state)", "fp16", {"DP_ACT": "fast"}),
    ("fp16 (default: fp16 operands, exp-based activations, fp32 state)", "fp16", {}),
    ("fp16 + folded weights hi+lo, MUFU.TANH", "fp16", {"DP_FOLD_SPLIT": "1", "DP_ACT": "fast"}),
    ("fp16 + folded weights hi+lo", "fp16", {"DP_FOLD_SPLIT": "1"}),
    ("fp16 + GRU layer 0 split", "fp16", {"DP_GRU_SPLIT": "1"}),
    ("fp16 + all GRU layers split", "fp16", {"DP_GRU_SPLIT": "7"}),
    ("fp16 + 16-bit state only, MUFU.TANH", "fp16", {"DP_GRU_STATE": "bf16", "DP_ACT": "fast"}),
    ("fp16 + 16-bit state only", "fp16", {"DP_GRU_STATE": "bf16"}),
    ("fp16x3 (all split, exp-based)", "fp16x3", {}),
]


def votes_of(pred):
    """test.py:71-82 over pred [nw, 148, 3] -> [nw, 146] (rows 2..147)."""
    a, b, c = pred[:, 0:146, 2].astype(np.int64), pred[:, 1:147, 1].astype(np.int64), pred[:, 2:148, 0].astype(np.int64)
    return np.where((a == b) | (a == c), a, np.where(b == c, b, np.minimum(a, np.minimum(b, c))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "model_error_attribution.json"))
    ap.add_argument("--only", default="", help="comma-separated substrings of configuration names")
    args = ap.parse_args()
    from deepore_b200 import model, modelio
    g4 = np.load(os.path.join(GOLD, "model_seed7.npz"))
    gc = np.load(os.path.join(GOLD, "bundled_calls_seed7.npz"))
    gi = np.load(os.path.join(GOLD, "bundled_na12878.npz"))
    img260 = torch.from_numpy(gi["img"]).cuda()
    img4 = torch.from_numpy(g4["img"]).cuda()
    have_logits = "logits" in gc.files
    ref_votes = votes_of(gc["pred"])
    assert np.array_equal(ref_votes, gc["calls"][:, 2:148])
    state = modelio.seeded_state(int(gc["seed"]))
    rows = []
    for name, precision, env in CONFIGS:
        if args.only and not any(s in name for s in args.only.split(",")):
            continue
        for k in SWITCHES:
            os.environ.pop(k, None)
        os.environ.update(env)
        m = model.from_state(state, "cuda", precision)
        lg4, _ = m.forward_windows(img4)
        lg, pred = m.forward_windows(img260)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(3):
            m.forward_windows(img260, want_logits=False)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 3
        d4 = lg4.cpu().numpy()[:, :148] - g4["logits"]
        pr = pred.cpu().numpy()[:, :148]
        r = {"config": name, "precision": precision, "env": env, "ms_per_260_windows": ms,
             "logits_maxabs_4win": float(np.abs(d4).max()), "logits_rms_4win": float(np.sqrt((d4 ** 2).mean())),
             "logit_scale_4win": float(np.abs(g4["logits"]).max())}
        diff = pr != gc["pred"]
        r["n_argmax"] = int(diff.size)
        r["n_argmax_diff"] = int(diff.sum())
        r["max_ref_margin_at_argmax_diff"] = float(gc["margin"][diff].max()) if diff.any() else 0.0
        v = votes_of(pr)
        r["n_votes"] = int(v.size)
        r["n_vote_diff"] = int((v != ref_votes).sum())
        if have_logits:
            d = lg.cpu().numpy()[:, :148] - gc["logits"]
            r["logits_maxabs_260win"] = float(np.abs(d).max())
            r["logits_rms_260win"] = float(np.sqrt((d ** 2).mean()))
            r["logit_scale_260win"] = float(np.abs(gc["logits"]).max())
        rows.append(r)
        print("%-62s max %.2e rms %.2e | argmax diff %5d / %d (margin <= %.1e) | vote diff %4d / %d | %.2f ms"
              % (name, r.get("logits_maxabs_260win", r["logits_maxabs_4win"]),
                 r.get("logits_rms_260win", r["logits_rms_4win"]), r["n_argmax_diff"], r["n_argmax"],
                 r["max_ref_margin_at_argmax_diff"], r["n_vote_diff"], r["n_votes"], ms), flush=True)
        del m
        torch.cuda.empty_cache()
    for k in SWITCHES:
        os.environ.pop(k, None)
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    json.dump(rows, open(args.out, "w"), indent=1)


if __name__ == "__main__":
    main()
