"""Synthetic workloads of SURVEY.md 8(d) / BASELINE.json configs 2,3,5: a random reference and
nanopore-like reads (lognormal lengths, substitution/insertion/deletion errors, soft clips), produced
directly in the SoA layout of `ReadBatch`.  Benchmark/test input generation only -- no reference code
corresponds to it (the reference ships no generator).
"""
from __future__ import annotations

from typing import List, Optional, Tuple

import numpy as np

from .reads import ReadBatch

_BASES = np.frombuffer(b"ACGT", dtype=np.uint8)
_CODE_OF = np.zeros(256, dtype=np.uint8)
for _i, _c in enumerate("=ACMGRSVTWYHKDBN"):
    _CODE_OF[ord(_c)] = _i


def random_reference(length: int, seed: int = 1, n_frac: float = 0.0) -> np.ndarray:
    rng = np.random.default_rng(seed)
    ref = _BASES[rng.integers(0, 4, size=length, dtype=np.int8)]
    if n_frac > 0:
        ref = ref.copy()
        ref[rng.random(length) < n_frac] = ord("N")
    return ref


def _geometric(rng, p, size):
    return rng.geometric(p, size=size).astype(np.int64)


def simulate_reads(ref: np.ndarray, ref_base: int = 0, coverage: float = 30.0, seed: int = 2,
                   mean_len: float = 9000.0, sigma: float = 0.5, min_len: int = 500, max_len: int = 100_000,
                   sub: float = 0.04, ins: float = 0.03, dele: float = 0.03, indel_p: float = 0.6,
                   clip_frac: float = 0.10, clip_max: int = 200, lowmapq_frac: float = 0.05,
                   region: Optional[Tuple[int, int]] = None, contig: str = "chr3",
                   read_n_frac: float = 0.0) -> ReadBatch:
    """Reads over ref (genomic [ref_base, ref_base+len(ref))).  If `region`=(lo,hi) is given, the full
    start/length draw is made (so the read set is a deterministic function of the seed) but only reads
    overlapping [lo,hi) are materialised."""
    rng = np.random.default_rng(seed)
    G = int(ref.shape[0])
    n_reads = max(1, int(round(coverage * G / mean_len / np.exp(sigma * sigma / 2.0))))
    lens = np.clip(np.exp(rng.normal(np.log(mean_len), sigma, size=n_reads)), min_len, max_len).astype(np.int64)
    lens = np.minimum(lens, G)
    starts = (rng.random(n_reads) * (G - lens + 1)).astype(np.int64)
    order = np.argsort(starts, kind="stable")
    starts, lens = starts[order], lens[order]
    seeds = rng.integers(0, 2 ** 62, size=n_reads)
    if region is not None:
        lo, hi = region[0] - ref_base, region[1] - ref_base
        keep = (starts < hi) & (starts + lens > lo)
        starts, lens, seeds = starts[keep], lens[keep], seeds[keep]

    pos, mapq, lseq = [], [], []
    cig_parts: List[np.ndarray] = []
    seq_parts: List[np.ndarray] = []
    qual_parts: List[np.ndarray] = []
    coff, soff, qoff = [0], [0], [0]
    p_event = ins + dele
    for s0, L, sd in zip(starts.tolist(), lens.tolist(), seeds.tolist()):
        r = np.random.default_rng(sd)
        # alternate match runs and indel events until L reference bases are consumed
        n_ev = max(4, int(L * p_event * 1.5) + 8)
        while True:
            mrun = _geometric(r, p_event, n_ev)                # >=1 matched bases between events
            ev_is_ins = r.random(n_ev) < (ins / p_event)
            ev_len = _geometric(r, indel_p, n_ev)
            ref_use = mrun + np.where(ev_is_ins, 0, ev_len)
            cs = np.cumsum(ref_use)
            if cs[-1] >= L:
                break
            n_ev *= 2
        k = int(np.searchsorted(cs, L, side="left"))          # events 0..k-1 complete; run k is cut
        consumed = int(cs[k - 1]) if k > 0 else 0
        last_m = L - consumed
        ops = np.empty(2 * k + 1, dtype=np.int64)
        ln = np.empty(2 * k + 1, dtype=np.int64)
        ops[0:2 * k:2] = 0
        ln[0:2 * k:2] = mrun[:k]
        ops[1:2 * k:2] = np.where(ev_is_ins[:k], 1, 2)
        ln[1:2 * k:2] = ev_len[:k]
        ops[2 * k] = 0
        ln[2 * k] = max(1, min(last_m, int(mrun[k])))
        # reference span actually used
        ref_span = int(np.sum(ln[(ops == 0) | (ops == 2)]))
        if s0 + ref_span > G:
            continue
        # read bases
        is_m = ops == 0
        is_i = ops == 1
        qlen = int(np.sum(ln[is_m | is_i]))
        seq = np.empty(qlen, dtype=np.uint8)
        q_off = np.concatenate([[0], np.cumsum(np.where(is_m | is_i, ln, 0))])[:-1]
        r_off = np.concatenate([[0], np.cumsum(np.where(is_m | (ops == 2), ln, 0))])[:-1]
        # matches: copy reference runs
        m_idx = np.flatnonzero(is_m)
        m_len = ln[m_idx]
        tot_m = int(m_len.sum())
        rep = np.repeat(np.arange(m_idx.size), m_len)
        within = np.arange(tot_m) - np.repeat(np.concatenate([[0], np.cumsum(m_len)])[:-1], m_len)
        qpos = q_off[m_idx][rep] + within
        rpos = s0 + r_off[m_idx][rep] + within
        seq[qpos] = ref[rpos]
        # substitutions
        sub_mask = r.random(tot_m) < sub
        nsub = int(sub_mask.sum())
        if nsub:
            cur = seq[qpos[sub_mask]]
            curi = np.searchsorted(_BASES, cur) % 4
            newi = (curi + r.integers(1, 4, size=nsub)) % 4
            seq[qpos[sub_mask]] = _BASES[newi]
        # inserted bases
        i_idx = np.flatnonzero(is_i)
        if i_idx.size:
            i_len = ln[i_idx]
            tot_i = int(i_len.sum())
            repi = np.repeat(np.arange(i_idx.size), i_len)
            withini = np.arange(tot_i) - np.repeat(np.concatenate([[0], np.cumsum(i_len)])[:-1], i_len)
            seq[q_off[i_idx][repi] + withini] = _BASES[r.integers(0, 4, size=tot_i)]
        if read_n_frac > 0:
            seq[r.random(qlen) < read_n_frac] = ord("N")
        # soft clips
        lead = trail = 0
        if r.random() < clip_frac:
            lead = int(r.integers(1, clip_max + 1))
        if r.random() < clip_frac:
            trail = int(r.integers(1, clip_max + 1))
        if lead or trail:
            seq = np.concatenate([_BASES[r.integers(0, 4, size=lead)], seq, _BASES[r.integers(0, 4, size=trail)]])
            if lead:
                ops = np.concatenate([[4], ops]); ln = np.concatenate([[lead], ln])
            if trail:
                ops = np.concatenate([ops, [4]]); ln = np.concatenate([ln, [trail]])
        qual = np.clip(np.rint(r.normal(12.0, 4.0, size=seq.shape[0])), 1, 40).astype(np.uint8)
        mq = 60 if r.random() >= lowmapq_frac else int(r.integers(0, 20))
        # pack
        codes = _CODE_OF[seq]
        if codes.shape[0] & 1:
            codes = np.concatenate([codes, np.zeros(1, np.uint8)])
        packed = (codes[0::2] << 4) | codes[1::2]
        pos.append(s0 + ref_base); mapq.append(mq); lseq.append(int(seq.shape[0]))
        cig_parts.append(((ln << 4) | ops).astype(np.uint32)); coff.append(coff[-1] + int(ops.shape[0]))
        seq_parts.append(packed); soff.append(soff[-1] + int(packed.shape[0]))
        qual_parts.append(qual); qoff.append(qoff[-1] + int(qual.shape[0]))

    def cat(parts, dt):
        return np.concatenate(parts).astype(dt) if parts else np.zeros(0, dt)

    return ReadBatch(contig=contig,
                     pos=np.asarray(pos, np.int32), flag=np.zeros(len(pos), np.uint16),
                     mapq=np.asarray(mapq, np.uint8), l_seq=np.asarray(lseq, np.int32),
                     cigar_off=np.asarray(coff, np.int64), cigar=cat(cig_parts, np.uint32),
                     seq_off=np.asarray(soff, np.int64), seq4=cat(seq_parts, np.uint8),
                     qual_off=np.asarray(qoff, np.int64), qual=cat(qual_parts, np.uint8))


def window_starts(region_lo: int, region_hi: int, stride: int = 100, flank: int = 100) -> np.ndarray:
    """queryStart of windows whose centres tile [region_lo, region_hi) at `stride`."""
    centres = np.arange(region_lo + flank, region_hi - flank - 1, stride, dtype=np.int64)
    return (centres - flank).astype(np.int32)


def bam_stream(rb, contig: str = "chr3", contig_len: int = 9_000_000) -> bytes:
    """Uncompressed BAM stream of a ReadBatch: magic, header with one reference, then one record per read in the layout of
    dep/htslib/htslib/sam.h:154-195 (name "r", no aux fields) -- the bytes a `samtools view -b` of these alignments holds."""
    import struct
    parts = [b"BAM\1", struct.pack("<i", 0), struct.pack("<i", 1), struct.pack("<i", len(contig) + 1),
             contig.encode() + b"\0", struct.pack("<i", contig_len)]
    cig = np.ascontiguousarray(rb.cigar, dtype="<u4")
    for i in range(rb.n_reads):
        c = cig[rb.cigar_off[i]:rb.cigar_off[i + 1]].tobytes()
        s = rb.seq4[rb.seq_off[i]:rb.seq_off[i + 1]].tobytes()
        q = rb.qual[rb.qual_off[i]:rb.qual_off[i + 1]].tobytes()
        body = struct.pack("<iiBBHHHiiii", 0, int(rb.pos[i]), 2, int(rb.mapq[i]), 0, len(c) // 4, int(rb.flag[i]),
                           int(rb.l_seq[i]), -1, -1, 0) + b"r\0" + c + s + q
        parts.append(struct.pack("<i", len(body)))
        parts.append(body)
    return b"".join(parts)


def bgzf_image(data: bytes, level: int = 6, chunk: int = 0xff00, threads: int = 0) -> np.ndarray:
    """BGZF file image of `data` as a uint8 array: blocks of `chunk` input bytes (htslib's BGZF_BLOCK_SIZE), 18-byte header
    with the 'BC' BSIZE sub-field, raw DEFLATE payload from zlib (the library htslib itself calls, dep/htslib/bgzf.c), CRC32 +
    ISIZE, the empty EOF block last.  zlib releases the GIL, so `threads` > 1 compresses blocks in parallel."""
    import struct
    import zlib

    def block(raw: bytes) -> bytes:
        co = zlib.compressobj(level, zlib.DEFLATED, -15)
        payload = co.compress(raw) + co.flush()
        bsize = 18 + len(payload) + 8
        if bsize > 65536:                                      # incompressible input: a stored block always fits
            co = zlib.compressobj(0, zlib.DEFLATED, -15)
            payload = co.compress(raw) + co.flush()
            bsize = 18 + len(payload) + 8
        head = struct.pack("<BBBBIBBHBBHH", 31, 139, 8, 4, 0, 0, 255, 6, 66, 67, 2, bsize - 1)
        return head + payload + struct.pack("<II", zlib.crc32(raw) & 0xffffffff, len(raw))

    pieces = [data[i:i + chunk] for i in range(0, len(data), chunk)]
    if threads and threads > 1 and len(pieces) > 1:
        from concurrent.futures import ThreadPoolExecutor
        with ThreadPoolExecutor(max_workers=threads) as ex:
            out = list(ex.map(block, pieces))
    else:
        out = [block(x) for x in pieces]
    out.append(block(b""))
    return np.frombuffer(b"".join(out), dtype=np.uint8)
