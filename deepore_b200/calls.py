"""Call set of a run: votes + image-row -> position map -> one record per scored row, written as TSV.

The reference stops at a confusion tensor (test.py:68-89, 109-111): it votes per image row of a window and never maps
the row back to the genome.  The CUDA path keeps that map (`rowpos`, `rowins` of dp_encode_windows: the reference
position of every image row and its index inside an insertion, Pileup.py:295-317), so the votes of dp_vote become
records (window, row, position, insert index, class).  Host-side numpy only: the arrays are the few bytes per row that
bench.py's e2e arm copies back anyway.

Windows overlap (window centres every `stride` < 201 bp), so a position can be scored by more than one window.  The
reference has no rule for that; `merge_records` keeps, per (position, insert index), the vote of the window whose centre
is nearest (ties: the earlier window), which is the window that sees the most context on both sides.
"""
from __future__ import annotations

from typing import Optional

import numpy as np

RECORD = np.dtype([("window", np.int32), ("row", np.int16), ("pos", np.int64), ("ins", np.int32), ("call", np.int8)])
HEADER = "#chrom\tpos\tins\tcall\twindow\trow\n"


def call_records(calls: np.ndarray, rowpos: np.ndarray, rowins: np.ndarray, window0: int = 0) -> np.ndarray:
    """calls int8 [Nw, H] (-1: row not scored, dp_vote), rowpos int32 [Nw, H] (0-based reference position of the row, -1:
    blank row), rowins int32 [Nw, H] (0: anchor row, k: k-th inserted base after it) -> RECORD array in (window, row)
    order.  `window0` is added to the window index (shards of one run)."""
    calls = np.asarray(calls)
    rowpos = np.asarray(rowpos)
    rowins = np.asarray(rowins)
    if calls.shape != rowpos.shape or calls.shape != rowins.shape or calls.ndim != 2:
        raise ValueError("calls, rowpos and rowins must be [n_windows, H] arrays of one shape")
    w, r = np.nonzero((calls >= 0) & (rowpos >= 0))
    rec = np.empty(w.shape[0], dtype=RECORD)
    rec["window"] = w + window0
    rec["row"] = r
    rec["pos"] = rowpos[w, r]
    rec["ins"] = rowins[w, r]
    rec["call"] = calls[w, r]
    return rec


def merge_records(rec: np.ndarray, win_qs: np.ndarray, flank: int = 100, window0: int = 0) -> np.ndarray:
    """One record per (pos, ins): of the windows that scored it, the one whose centre (win_qs + flank) is nearest; ties go to
    the earlier window.  Output sorted by (pos, ins)."""
    if rec.shape[0] == 0:
        return rec.copy()
    centre = np.asarray(win_qs, dtype=np.int64)[rec["window"] - window0] + flank
    dist = np.abs(rec["pos"] - centre)
    order = np.lexsort((rec["window"], dist, rec["ins"], rec["pos"]))
    s = rec[order]
    first = np.ones(s.shape[0], dtype=bool)
    first[1:] = (s["pos"][1:] != s["pos"][:-1]) | (s["ins"][1:] != s["ins"][:-1])
    return s[first]


def write_tsv(path: str, chrom: str, rec: np.ndarray, append: bool = False, only_class: Optional[int] = None) -> int:
    """chrom, 1-based position, insert index, class, window, row -- one line per record; returns the number written."""
    if only_class is not None:
        rec = rec[rec["call"] == only_class]
    with open(path, "a" if append else "w") as f:
        if not append:
            f.write(HEADER)
        for x in rec:
            f.write("%s\t%d\t%d\t%d\t%d\t%d\n" % (chrom, int(x["pos"]) + 1, int(x["ins"]), int(x["call"]), int(x["window"]),
                                                  int(x["row"])))
    return int(rec.shape[0])


def read_tsv(path: str):
    """-> (list of chromosome names, RECORD array); inverse of write_tsv."""
    chroms, rows = [], []
    with open(path) as f:
        for ln in f:
            if not ln.strip() or ln.startswith("#"):
                continue
            c, pos, ins, call, win, row = ln.rstrip("\n").split("\t")
            chroms.append(c)
            rows.append((int(win), int(row), int(pos) - 1, int(ins), int(call)))
    return chroms, np.array(rows, dtype=RECORD) if rows else np.empty(0, dtype=RECORD)
