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


# Classes as the reference's labels define them (pileupGenerator.py:23-32 getClassForGenotype; '0' is the pad label of
# Pileup.py:374-378): 0 = no label, 1 = homozygous reference, 2 = heterozygous, 3 = homozygous alternate.  The reference
# model is built with class_n = 3 (model.py:60), so class 3 never comes out of it; the map is a parameter for that reason.
CLASS_GENOTYPES = {1: "0/0", 2: "0/1", 3: "1/1"}


def write_vcf(path: str, chrom: str, rec: np.ndarray, ref: bytes, ref_base: int = 0, sample: str = "SAMPLE",
              class_genotypes: Optional[dict] = None, emit_hom_ref: bool = False) -> int:
    """Sites-only VCF 4.2 of a merged call set: one line per record whose class maps to a genotype (hom-ref lines only
    with `emit_hom_ref`).  The model predicts a zygosity class per image row, not an allele (model.py:154-155,
    test.py:68-89), so ALT is the symbolic <NON_REF>; rows inside an insertion (ins = k > 0) are written at their anchor
    position with ALT <INS> and INFO INSROW=k.  `ref`: upper-case reference bytes, `ref[p - ref_base]` = base at 0-based p.
    Records must be sorted by (pos, ins) (merge_records' order); returns the number of lines written."""
    gmap = CLASS_GENOTYPES if class_genotypes is None else class_genotypes
    if rec.shape[0] > 1:
        k = rec["pos"].astype(np.int64) * (1 << 20) + rec["ins"]
        if np.any(k[1:] < k[:-1]):
            raise ValueError("records must be sorted by (pos, ins); run merge_records first")
    n = 0
    with open(path, "w") as f:
        f.write("##fileformat=VCFv4.2\n##source=deepore_b200\n")
        f.write('##ALT=<ID=NON_REF,Description="Any alternate allele (the model scores zygosity, not alleles)">\n')
        f.write('##ALT=<ID=INS,Description="Row inside an insertion after POS">\n')
        f.write('##INFO=<ID=INSROW,Number=1,Type=Integer,Description="Index of the inserted base the row stands for">\n')
        f.write('##INFO=<ID=CLASS,Number=1,Type=Integer,Description="Class voted for the row (test.py:68-89)">\n')
        f.write('##FORMAT=<ID=GT,Number=1,Type=String,Description="Genotype of the voted class">\n')
        f.write("#CHROM\tPOS\tID\tREF\tALT\tQUAL\tFILTER\tINFO\tFORMAT\t%s\n" % sample)
        for x in rec:
            c = int(x["call"])
            gt = gmap.get(c)
            if gt is None or (not emit_hom_ref and set(gt.replace("|", "/").split("/")) == {"0"}):
                continue
            p = int(x["pos"])
            i = p - ref_base
            if i < 0 or i >= len(ref):
                raise IndexError("position %d outside the reference slice [%d, %d)" % (p, ref_base, ref_base + len(ref)))
            rb = chr(ref[i]) if isinstance(ref, (bytes, bytearray, np.ndarray)) else ref[i]
            ins = int(x["ins"])
            alt, info = ("<INS>", "CLASS=%d;INSROW=%d" % (c, ins)) if ins > 0 else ("<NON_REF>", "CLASS=%d" % c)
            f.write("%s\t%d\t.\t%s\t%s\t.\t.\t%s\tGT\t%s\n" % (chrom, p + 1, rb, alt, info, gt))
            n += 1
    return n
