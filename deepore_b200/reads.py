"""Host-side read containers: BAM/FASTA access and the structure-of-arrays read batch that the CUDA
encoder consumes (include/deepore_b200.h `dp_reads`).

Replaces the data-access side of the reference encoder -- `pysam.AlignmentFile.fetch` and
`pyfaidx.Fasta.get_seq` at /root/reference/src/utils/Pileup.py:36,41,414-415 -- with one pass that
inflates the BGZF stream and lays every record of a contig out as flat arrays in file order.  Record
layout per dep/htslib/htslib/sam.h:154-195; region overlap rule per dep/htslib/hts.c:2242 and
bam_endpos per dep/htslib/sam.c:347-353 (applied again, exactly, inside the kernel).
"""
from __future__ import annotations

import gzip
import os
import struct
from dataclasses import dataclass, field
from typing import Dict, Iterable, List, Optional, Sequence, Tuple

import numpy as np

NT16 = "=ACMGRSVTWYHKDBN"
_NT16_CODE = {c: i for i, c in enumerate(NT16)}
CIGAR_OPS = "MIDNSHP=XB"
_CONSUMES_REF = np.array([1, 0, 1, 1, 0, 0, 0, 1, 1, 0, 0, 0, 0, 0, 0, 0], dtype=np.int64)
BAM_FUNMAP = 0x4


def pack_seq(seq: str) -> bytes:
    """ASCII bases -> BAM 4-bit packing (high nibble first)."""
    codes = [_NT16_CODE[c] for c in seq]
    if len(codes) & 1:
        codes.append(0)
    return bytes((codes[i] << 4) | codes[i + 1] for i in range(0, len(codes), 2))


@dataclass
class ReadBatch:
    """All records of one contig, file order, SoA.  Arrays are numpy on the host; `to_device()` gives the
    torch mirror whose pointers go into `dp_reads`."""
    contig: str
    pos: np.ndarray        # int32 [n]
    flag: np.ndarray       # uint16 [n]
    mapq: np.ndarray       # uint8 [n]
    l_seq: np.ndarray      # int32 [n]
    cigar_off: np.ndarray  # int64 [n+1]
    cigar: np.ndarray      # uint32 [ncig]
    seq_off: np.ndarray    # int64 [n+1] (bytes)
    seq4: np.ndarray       # uint8
    qual_off: np.ndarray   # int64 [n+1]
    qual: np.ndarray       # uint8
    _endpos: Optional[np.ndarray] = field(default=None, repr=False)
    _prefmax_end: Optional[np.ndarray] = field(default=None, repr=False)
    _dev: Dict[str, object] = field(default_factory=dict, repr=False)

    @property
    def n_reads(self) -> int:
        return int(self.pos.shape[0])

    @property
    def n_cigar(self) -> int:
        return int(self.cigar.shape[0])

    # ------------------------------------------------------------------ construction
    @staticmethod
    def from_records(records: Iterable[dict], contig: str = "chr3") -> "ReadBatch":
        """records: dicts with pos, mapq, cigar [(op,len)...] or "10M2I..", seq (str), qual (bytes/list),
        optional flag.  Kept in the given order (callers pass coordinate-sorted records, like a BAM)."""
        pos, flag, mapq, lseq = [], [], [], []
        coff, soff, qoff = [0], [0], [0]
        cig: List[int] = []
        seqb: List[bytes] = []
        qualb: List[bytes] = []
        for r in records:
            c = r["cigar"]
            if isinstance(c, str):
                c = parse_cigar_string(c)
            seq = r["seq"]
            q = bytes(r["qual"])
            assert len(q) == len(seq), "qual/seq length mismatch"
            pos.append(r["pos"]); flag.append(r.get("flag", 0)); mapq.append(r["mapq"]); lseq.append(len(seq))
            cig.extend((n << 4) | op for op, n in c)
            coff.append(len(cig))
            sb = pack_seq(seq)
            seqb.append(sb); soff.append(soff[-1] + len(sb))
            qualb.append(q); qoff.append(qoff[-1] + len(q))
        return ReadBatch(
            contig=contig,
            pos=np.asarray(pos, dtype=np.int32), flag=np.asarray(flag, dtype=np.uint16),
            mapq=np.asarray(mapq, dtype=np.uint8), l_seq=np.asarray(lseq, dtype=np.int32),
            cigar_off=np.asarray(coff, dtype=np.int64), cigar=np.asarray(cig, dtype=np.uint32),
            seq_off=np.asarray(soff, dtype=np.int64),
            seq4=np.frombuffer(b"".join(seqb), dtype=np.uint8).copy() if seqb else np.zeros(0, np.uint8),
            qual_off=np.asarray(qoff, dtype=np.int64),
            qual=np.frombuffer(b"".join(qualb), dtype=np.uint8).copy() if qualb else np.zeros(0, np.uint8))

    @staticmethod
    def from_bam(path: str, contig: str) -> "ReadBatch":
        return BamFile(path).read_contig(contig)

    # ------------------------------------------------------------------ derived host indices
    @property
    def endpos(self) -> np.ndarray:
        """bam_endpos (dep/htslib/sam.c:347-353) for every record."""
        if self._endpos is None:
            oplen = (self.cigar >> 4).astype(np.int64) * _CONSUMES_REF[self.cigar & 0xF]
            cs = np.concatenate([[0], np.cumsum(oplen)])
            rlen = cs[self.cigar_off[1:]] - cs[self.cigar_off[:-1]]
            ncig = np.diff(self.cigar_off)
            mapped = ((self.flag & BAM_FUNMAP) == 0) & (ncig > 0)
            self._endpos = np.where(mapped, self.pos.astype(np.int64) + rlen, self.pos.astype(np.int64) + 1)
        return self._endpos

    def window_ranges(self, win_qs: np.ndarray, flank: int) -> Tuple[np.ndarray, np.ndarray]:
        """Candidate range [lo,hi) of records per window: a superset of `fetch(start=qs, end=qe)`
        (pos < qe && endpos > qs) for coordinate-sorted records; the encoder applies the exact rule."""
        win_qs = np.asarray(win_qs, dtype=np.int64)
        qe = win_qs + (2 * flank + 1)
        if self.n_reads and np.any(np.diff(self.pos) < 0):
            raise ValueError("records must be coordinate-sorted (as pysam.fetch requires an index)")
        if self._prefmax_end is None:
            self._prefmax_end = np.maximum.accumulate(self.endpos) if self.n_reads else np.zeros(0, np.int64)
        lo = np.searchsorted(self._prefmax_end, win_qs, side="right")
        hi = np.searchsorted(self.pos, qe, side="left")
        hi = np.maximum(hi, lo)
        return lo.astype(np.int32), hi.astype(np.int32)

    # ------------------------------------------------------------------ device mirror
    def to_device(self, device="cuda", stream=None):
        """Upload once (pinned staging + non_blocking copies) and cache the torch tensors."""
        import torch
        key = str(device)
        if key not in self._dev:
            d = {}
            for name in ("pos", "flag", "mapq", "l_seq", "cigar_off", "cigar", "seq_off", "seq4", "qual_off", "qual"):
                a = getattr(self, name)
                if a.dtype == np.uint16:
                    t = torch.from_numpy(a.view(np.int16))
                elif a.dtype == np.uint32:
                    t = torch.from_numpy(a.view(np.int32))
                else:
                    t = torch.from_numpy(a)
                if t.numel() == 0:
                    t = torch.zeros(16, dtype=t.dtype)
                if torch.device(device).type == "cuda":
                    t = t.pin_memory().to(device, non_blocking=True)
                d[name] = t
            self._dev[key] = d
        return self._dev[key]

    def slice(self, r0: int, r1: int) -> "ReadBatch":
        """Records [r0, r1) as a self-contained batch (offsets rebased) -- what a host reader hands over for
        one genomic interval."""
        r0 = max(0, int(r0)); r1 = max(r0, min(int(r1), self.n_reads))
        c0, c1 = int(self.cigar_off[r0]), int(self.cigar_off[r1])
        s0, s1 = int(self.seq_off[r0]), int(self.seq_off[r1])
        q0, q1 = int(self.qual_off[r0]), int(self.qual_off[r1])
        return ReadBatch(contig=self.contig, pos=self.pos[r0:r1].copy(), flag=self.flag[r0:r1].copy(),
                         mapq=self.mapq[r0:r1].copy(), l_seq=self.l_seq[r0:r1].copy(),
                         cigar_off=self.cigar_off[r0:r1 + 1] - c0, cigar=self.cigar[c0:c1].copy(),
                         seq_off=self.seq_off[r0:r1 + 1] - s0, seq4=self.seq4[s0:s1].copy(),
                         qual_off=self.qual_off[r0:r1 + 1] - q0, qual=self.qual[q0:q1].copy())

    def nbytes(self) -> int:
        return sum(getattr(self, n).nbytes for n in
                   ("pos", "flag", "mapq", "l_seq", "cigar_off", "cigar", "seq_off", "seq4", "qual_off", "qual"))


def parse_cigar_string(s: str) -> List[Tuple[int, int]]:
    out, num = [], ""
    for ch in s:
        if ch.isdigit():
            num += ch
        else:
            out.append((CIGAR_OPS.index(ch), int(num)))
            num = ""
    return out


class BamFile:
    """Minimal sequential BAM reader (BGZF members inflate with zlib; no index needed: one pass, whole
    file).  Host code; the north star keeps BAM decompression/parsing on the host."""

    def __init__(self, path: str):
        self.path = path
        with gzip.open(path, "rb") as fh:
            self._data = fh.read()
        d = self._data
        if d[:4] != b"BAM\1":
            raise ValueError("%s: not a BAM file" % path)
        l_text = struct.unpack_from("<i", d, 4)[0]
        off = 8 + l_text
        n_ref = struct.unpack_from("<i", d, off)[0]
        off += 4
        self.references: List[str] = []
        self.lengths: List[int] = []
        for _ in range(n_ref):
            l_name = struct.unpack_from("<i", d, off)[0]
            self.references.append(d[off + 4:off + 4 + l_name - 1].decode("ascii"))
            self.lengths.append(struct.unpack_from("<i", d, off + 4 + l_name)[0])
            off += 8 + l_name
        self._first_record = off

    def read_contig(self, contig: str) -> ReadBatch:
        tid = self.references.index(contig)
        d = self._data
        mv = memoryview(d)
        off, n = self._first_record, len(d)
        pos, flag, mapq, lseq = [], [], [], []
        coff, soff, qoff = [0], [0], [0]
        cig_parts, seq_parts, qual_parts = [], [], []
        unpack = struct.Struct("<iiiBBHHHi").unpack_from
        while off < n:
            block_size, ref_id, p, l_name, mq, _bin, n_cig, fl, l_seq = unpack(d, off)
            if ref_id == tid:
                q = off + 36 + l_name
                pos.append(p); flag.append(fl); mapq.append(mq); lseq.append(l_seq)
                cig_parts.append(mv[q:q + 4 * n_cig]); coff.append(coff[-1] + n_cig)
                q += 4 * n_cig
                nb = (l_seq + 1) >> 1
                seq_parts.append(mv[q:q + nb]); soff.append(soff[-1] + nb)
                q += nb
                qual_parts.append(mv[q:q + l_seq]); qoff.append(qoff[-1] + l_seq)
            off += 4 + block_size
        cig = np.frombuffer(b"".join(cig_parts), dtype="<u4").astype(np.uint32)
        return ReadBatch(
            contig=contig,
            pos=np.asarray(pos, dtype=np.int32), flag=np.asarray(flag, dtype=np.uint16),
            mapq=np.asarray(mapq, dtype=np.uint8), l_seq=np.asarray(lseq, dtype=np.int32),
            cigar_off=np.asarray(coff, dtype=np.int64), cigar=cig,
            seq_off=np.asarray(soff, dtype=np.int64),
            seq4=np.frombuffer(b"".join(seq_parts), dtype=np.uint8).copy(),
            qual_off=np.asarray(qoff, dtype=np.int64),
            qual=np.frombuffer(b"".join(qual_parts), dtype=np.uint8).copy())


class FastaFile:
    """Plain FASTA access with the pyfaidx call convention the reference uses
    (`get_seq(name, start, end)`: 1-based inclusive, upper-cased; Pileup.py:41,415)."""

    def __init__(self, path: str):
        self.path = path
        self._seqs: Dict[str, np.ndarray] = {}
        name, chunks = None, []
        with open(path, "rb") as fh:
            for line in fh:
                if line.startswith(b">"):
                    if name is not None:
                        self._seqs[name] = self._finish(chunks)
                    name, chunks = line[1:].split()[0].decode("ascii"), []
                else:
                    chunks.append(line.strip())
        if name is not None:
            self._seqs[name] = self._finish(chunks)

    @staticmethod
    def _finish(chunks) -> np.ndarray:
        a = np.frombuffer(b"".join(chunks), dtype=np.uint8).copy()
        lower = (a >= ord("a")) & (a <= ord("z"))
        a[lower] -= 32
        return a

    def contig(self, name: str) -> np.ndarray:
        return self._seqs[name]

    def get_seq(self, name: str, start: int, end: int) -> str:
        return self._seqs[name][start - 1:end].tobytes().decode("ascii")

    def fetch(self, name: str, start0: int, end0: int) -> np.ndarray:
        """0-based half-open slice as uint8."""
        return self._seqs[name][start0:end0]
