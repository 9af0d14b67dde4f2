"""Interval shards: the unit of work of the B200 path.

The reference encodes one window per call and re-opens the BAM for each (pileupGenerator.py:111-112 ->
Pileup.py:413-451).  Here the unit is a genomic INTERVAL: every window whose centre lies in it, the reads
that overlap those windows (one contiguous run of a coordinate-sorted BAM) and the reference slice under
them, packed so that one shard == a handful of host->device copies and one kernel launch.  Shards are
independent, which is also how the path scales over GPUs: rank r takes shards r, r+N, ...
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Optional

import numpy as np

from .reads import ReadBatch, _CONSUMES_REF


@dataclass
class IntervalShard:
    reads: ReadBatch          # records [r0, r1) of the contig, offsets rebased
    ref: np.ndarray           # uint8 reference bases for [ref_base, ref_base+len)
    ref_base: int
    win_qs: np.ndarray        # int32 [Nw] queryStart of each window
    win_lo: np.ndarray        # int32 [Nw] candidate read range (indices into `reads`)
    win_hi: np.ndarray
    flank: int = 100

    @property
    def n_windows(self) -> int:
        return int(self.win_qs.shape[0])

    def input_nbytes(self) -> int:
        return self.reads.nbytes() + self.ref.nbytes + self.win_qs.nbytes + self.win_lo.nbytes + self.win_hi.nbytes


def make_shard(rb: ReadBatch, ref: np.ndarray, ref_base: int, win_qs: np.ndarray, flank: int = 100) -> IntervalShard:
    """Cut the shard of `win_qs` (ascending) out of a contig-wide batch."""
    win_qs = np.ascontiguousarray(win_qs, dtype=np.int32)
    L = 2 * flank + 1
    lo, hi = rb.window_ranges(win_qs, flank)
    if win_qs.shape[0] == 0:
        return IntervalShard(rb.slice(0, 0), ref[:0].copy(), ref_base, win_qs, lo, hi, flank)
    r0, r1 = int(lo.min()), int(hi.max())
    sub = rb.slice(r0, r1)
    g0 = max(int(win_qs.min()), ref_base)
    g1 = min(int(win_qs.max()) + L + 1, ref_base + int(ref.shape[0]))
    return IntervalShard(sub, np.ascontiguousarray(ref[g0 - ref_base:g1 - ref_base]), g0, win_qs,
                         (lo - r0).astype(np.int32), (hi - r0).astype(np.int32), flank)


def algorithmic_bytes(sh: IntervalShard, W: int = 200, H: int = 150, mapq_cut: int = 20) -> float:
    """SURVEY.md 8(d) encoder roofline numerator, summed over the shard's windows:
    H*W*4 (RGBA write) + (L+1) (reference) + sum over fetched reads of
    [in-window bases * 1.5 (4-bit base + quality byte) + in-window CIGAR ops * 4 + 16]."""
    rb = sh.reads
    L = 2 * sh.flank + 1
    nw = sh.n_windows
    total = float(nw) * (H * W * 4 + L + 1)
    if rb.n_reads == 0 or nw == 0:
        return total
    pos = rb.pos.astype(np.int64)
    end = rb.endpos.astype(np.int64)
    span = np.maximum(end - pos, 1)
    nops = np.diff(rb.cigar_off).astype(np.float64)
    dens = nops / span                                   # CIGAR ops per reference base of the read
    ok = rb.mapq >= mapq_cut
    # windows are few thousand, candidate ranges ~100 reads: vectorise per window over its range
    for w in range(nw):
        qs = int(sh.win_qs[w]); qe = qs + L
        a, b = int(sh.win_lo[w]), int(sh.win_hi[w])
        if b <= a:
            continue
        p, e = pos[a:b], end[a:b]
        ov = np.minimum(e, qe) - np.maximum(p, qs)
        m = (ov > 0) & ok[a:b]
        if not m.any():
            continue
        ovm = ov[m].astype(np.float64)
        total += float(np.sum(ovm * 1.5 + ovm * dens[a:b][m] * 4.0 + 16.0))
    return total


class SyntheticWorkload:
    """BASELINE.json configs[1] / SURVEY.md 8(d) config 2: 9 Mb i.i.d. reference (seed 1), 30x
    nanopore-like reads (seed 2), window centres every `stride` bp.  Only the part of the read set that the
    requested shards touch is materialised (the draw of starts/lengths is always the full one, so the read
    set is a function of the seeds alone)."""

    def __init__(self, ref_len: int = 9_000_000, coverage: float = 30.0, stride: int = 100, flank: int = 100,
                 ref_seed: int = 1, read_seed: int = 2):
        from . import synth
        self.synth = synth
        self.ref_len, self.coverage, self.stride, self.flank = ref_len, coverage, stride, flank
        self.read_seed = read_seed
        self.ref = synth.random_reference(ref_len, seed=ref_seed)
        self.all_win_qs = synth.window_starts(0, ref_len, stride=stride, flank=flank)

    @property
    def n_windows(self) -> int:
        return int(self.all_win_qs.shape[0])

    def shards(self, first_window: int, windows_per_shard: int, n_shards: int) -> List[IntervalShard]:
        nwin = self.n_windows
        w0 = first_window % nwin
        w1 = min(nwin, w0 + windows_per_shard * n_shards)
        wq_all = self.all_win_qs[w0:w1]
        L = 2 * self.flank + 1
        region = (int(wq_all[0]), int(wq_all[-1]) + L + 1)
        rb = self.synth.simulate_reads(self.ref, coverage=self.coverage, seed=self.read_seed, region=region)
        out = []
        for i in range(n_shards):
            wq = wq_all[i * windows_per_shard:(i + 1) * windows_per_shard]
            if wq.shape[0] == 0:
                break
            out.append(make_shard(rb, self.ref, 0, wq, self.flank))
        return out
