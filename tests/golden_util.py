"""Helpers shared by the tests: load a tests/golden/*.npz fixture as (ReadBatch, ref, ref_base, wins, img...)."""
import os
from types import SimpleNamespace

import numpy as np

from deepore_b200.reads import ReadBatch

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
FIXTURES = ["quirks", "deep", "fuzz", "nanopore_small", "bundled_na12878"]


def load(name):
    z = np.load(os.path.join(GOLD, name + ".npz"))
    rb = ReadBatch(contig="chr3", **{k: z[k] for k in
                                    ("pos", "flag", "mapq", "l_seq", "cigar_off", "cigar", "seq_off", "seq4",
                                     "qual_off", "qual")})
    extra = {k: z[k] for k in z.files}
    return SimpleNamespace(reads=rb, ref=z["ref"], ref_base=int(z["ref_base"]), win_qs=z["win_qs"],
                           img=z["img"], status=z["status"], labels=z["labels"], z=extra)


def trained_state(path=None):
    """modelio.seeded_state(seed) with the small tensors that oracle/make_trained_golden.py's SGD steps changed (stored under
    `state/<name>` in tests/golden/bundled_calls_trained.npz) put in place -> (state, the npz)."""
    from deepore_b200 import modelio
    z = np.load(path or os.path.join(GOLD, "bundled_calls_trained.npz"))
    st = dict(modelio.seeded_state(int(z["seed"])))
    for k in z.files:
        if k.startswith("state/"):
            name = k[len("state/"):]
            assert name in st and st[name].shape == z[k].shape, name
            st[name] = z[k]
    return st, z


def write_bam(path, rb, contig="chr3", contig_len=9_000_000):
    """Uncompressed-member BAM of a ReadBatch (record layout of dep/htslib/htslib/sam.h:154-195; names 'r', no aux fields):
    enough for `reads.BamFile`, which inflates the whole stream with gzip."""
    import gzip
    import struct
    parts = [b"BAM\1", struct.pack("<i", 0), struct.pack("<i", 1),
             struct.pack("<i", len(contig) + 1), contig.encode() + b"\0", struct.pack("<i", contig_len)]
    cig = rb.cigar.astype("<u4")
    for i in range(rb.n_reads):
        c = cig[rb.cigar_off[i]:rb.cigar_off[i + 1]].tobytes()
        s = rb.seq4[rb.seq_off[i]:rb.seq_off[i + 1]].tobytes()
        q = rb.qual[rb.qual_off[i]:rb.qual_off[i + 1]].tobytes()
        body = struct.pack("<iiBBHHHiiii", 0, int(rb.pos[i]), 2, int(rb.mapq[i]), 0, len(c) // 4, int(rb.flag[i]),
                           int(rb.l_seq[i]), -1, -1, 0) + b"r\0" + c + s + q
        parts.append(struct.pack("<i", len(body)))
        parts.append(body)
    with gzip.open(path, "wb", compresslevel=1) as fh:
        fh.write(b"".join(parts))


def write_fasta(path, ref, ref_base, contig="chr3", width=60):
    """FASTA whose contig is `ref` placed at 0-based `ref_base`, N before it."""
    s = b"N" * int(ref_base) + bytes(ref)
    with open(path, "wb") as fh:
        fh.write(b">" + contig.encode() + b"\n")
        for i in range(0, len(s), width):
            fh.write(s[i:i + width] + b"\n")


def bgzf_compress(data: bytes, level: int = 6, chunk: int = 0xff00, strategy: int = 0, eof: bool = True) -> bytes:
    """BGZF file image of `data` (block layout of dep/htslib/bgzf.c: 18-byte gzip header with the 'BC' BSIZE sub-field,
    raw DEFLATE payload, CRC32 + ISIZE; htslib's empty EOF block last).  Payloads come from zlib, the library htslib
    itself calls; `strategy` = zlib.Z_FIXED gives fixed-Huffman blocks, level 0 stored ones."""
    import struct
    import zlib

    def block(raw: bytes) -> bytes:
        co = zlib.compressobj(level, zlib.DEFLATED, -15, 9, strategy)
        payload = co.compress(raw) + co.flush()
        bsize = 18 + len(payload) + 8
        assert bsize <= 65536, "block does not fit BSIZE; lower `chunk`"
        head = struct.pack("<BBBBIBBHBBHH", 31, 139, 8, 4, 0, 0, 255, 6, 66, 67, 2, bsize - 1)
        return head + payload + struct.pack("<II", zlib.crc32(raw) & 0xffffffff, len(raw))

    out = [block(data[i:i + chunk]) for i in range(0, len(data), chunk)]
    if eof:
        out.append(block(b""))
    return b"".join(out)


def bam_bytes(rb, contig="chr3", contig_len=9_000_000, other_contig_every=0, names=("r",)):
    """Uncompressed BAM stream of a ReadBatch (same record layout as write_bam); with `other_contig_every` = k every k-th
    record is duplicated onto a second contig 'chrOther' (tid 1), which a contig filter has to drop; read names cycle
    through `names` so that records have different name lengths."""
    import struct
    parts = [b"BAM\1", struct.pack("<i", 4), b"@HD\n", struct.pack("<i", 2),
             struct.pack("<i", len(contig) + 1), contig.encode() + b"\0", struct.pack("<i", contig_len),
             struct.pack("<i", 9), b"chrOther\0", struct.pack("<i", 1000)]
    cig = rb.cigar.astype("<u4")
    for i in range(rb.n_reads):
        c = cig[rb.cigar_off[i]:rb.cigar_off[i + 1]].tobytes()
        s = rb.seq4[rb.seq_off[i]:rb.seq_off[i + 1]].tobytes()
        q = rb.qual[rb.qual_off[i]:rb.qual_off[i + 1]].tobytes()
        nm = names[i % len(names)].encode() + b"\0"
        for tid in ([0, 1] if other_contig_every and i % other_contig_every == 0 else [0]):
            body = struct.pack("<iiBBHHHiiii", tid, int(rb.pos[i]), len(nm), int(rb.mapq[i]), 0, len(c) // 4,
                               int(rb.flag[i]), int(rb.l_seq[i]), -1, -1, 0) + nm + c + s + q + b"NMC\x01"
            parts.append(struct.pack("<i", len(body)))
            parts.append(body)
    return b"".join(parts)
