"""TEST INFRASTRUCTURE ONLY -- stand-in for the absent third-party package `pysam`.

The reference's encoder (`/root/reference/src/utils/Pileup.py:1`) imports pysam, which is not
installed here and cannot be (no network).  This module supplies just the attributes the reference
touches (Pileup.py:36,39,40,160,183-185,193,197,414), with semantics taken from the vendored
htslib sources the real pysam wraps:

* BAM record layout ............ dep/htslib/htslib/sam.h:154-195
* region overlap rule .......... dep/htslib/hts.c:2242  (pos < end && endpos > beg, file order)
* bam_endpos ................... dep/htslib/sam.c:347-353
* 4-bit base alphabet .......... dep/htslib/hts.c:82 ("=ACMGRSVTWYHKDBN")

It is a plain sequential reader (no .bai use): every fetch() scans the cached record list.
Never imported by the product package.
"""
import gzip
import struct
from array import array

_NT16 = "=ACMGRSVTWYHKDBN"
# ops that consume reference: M D N = X  (sam.h:65-109 bam_cigar_type)
_CONSUMES_REF = (1, 0, 1, 1, 0, 0, 0, 1, 1, 0)


class AlignedSegment(object):
    __slots__ = ("reference_id", "reference_start", "mapping_quality", "flag", "cigartuples",
                 "_seq", "_qual", "query_name", "_aux", "reference_end_htslib")

    @property
    def query_sequence(self):
        return self._seq

    @property
    def query_qualities(self):
        # pysam returns None when the first quality byte is 0xff, else array('B') of the FULL query
        if self._qual is None:
            return None
        return array('B', self._qual)

    @property
    def query_alignment_start(self):
        start = 0
        if not self.cigartuples:
            return 0
        for op, n in self.cigartuples:
            if op == 5:          # H: allowed only at the ends
                continue
            if op == 4:
                start += n
            else:
                break
        return start

    @property
    def query_alignment_end(self):
        end = len(self._seq) if self._seq is not None else 0
        if not self.cigartuples or end == 0:
            return end
        for op, n in reversed(self.cigartuples):
            if op == 5:
                continue
            if op == 4:
                end -= n
            else:
                break
        return end

    @property
    def query_alignment_sequence(self):
        if self._seq is None:
            return None
        return self._seq[self.query_alignment_start:self.query_alignment_end]

    def get_reference_positions(self, full_length=False):
        out = []
        pos = self.reference_start
        if not self.cigartuples:
            return out
        for op, n in self.cigartuples:
            if op in (0, 7, 8):
                out.extend(range(pos, pos + n))
                pos += n
            elif op in (2, 3):
                pos += n
        return out

    def get_tag(self, tag):
        return self._aux[tag]

    def has_tag(self, tag):
        return tag in self._aux


def _parse_aux(buf):
    aux = {}
    i = 0
    n = len(buf)
    while i + 3 <= n:
        tag = buf[i:i + 2].decode("ascii")
        t = chr(buf[i + 2])
        i += 3
        if t == 'A':
            aux[tag] = chr(buf[i]); i += 1
        elif t in 'cC':
            aux[tag] = struct.unpack_from('<b' if t == 'c' else '<B', buf, i)[0]; i += 1
        elif t in 'sS':
            aux[tag] = struct.unpack_from('<h' if t == 's' else '<H', buf, i)[0]; i += 2
        elif t in 'iI':
            aux[tag] = struct.unpack_from('<i' if t == 'i' else '<I', buf, i)[0]; i += 4
        elif t == 'f':
            aux[tag] = struct.unpack_from('<f', buf, i)[0]; i += 4
        elif t in 'ZH':
            j = buf.index(b'\0', i)
            aux[tag] = buf[i:j].decode("ascii"); i = j + 1
        elif t == 'B':
            st = chr(buf[i]); cnt = struct.unpack_from('<I', buf, i + 1)[0]; i += 5
            sz = {'c': 1, 'C': 1, 's': 2, 'S': 2, 'i': 4, 'I': 4, 'f': 4}[st]
            i += sz * cnt
        else:
            raise ValueError("bad aux type %r" % t)
    return aux


def make_record(pos, mapq, cigartuples, seq, qual, flag=0, reference_id=0, name="r"):
    """Build an in-memory record (for synthetic vectors run through the unmodified reference)."""
    r = AlignedSegment()
    r.reference_id = reference_id
    r.reference_start = pos
    r.mapping_quality = mapq
    r.flag = flag
    r.query_name = name
    r.cigartuples = list(cigartuples) if cigartuples else None
    r._seq = seq if seq else None
    r._qual = bytes(qual) if seq else None
    r._aux = {}
    if not (flag & 4) and r.cigartuples:
        r.reference_end_htslib = pos + sum(ln for op, ln in r.cigartuples if _CONSUMES_REF[op])
    else:
        r.reference_end_htslib = pos + 1
    return r


def register_memory_file(path, references, lengths, records):
    """Make AlignmentFile(path) serve `records` (list of make_record results, file order)."""
    AlignmentFile._cache[path] = (list(references), list(lengths), list(records))


class AlignmentFile(object):
    _cache = {}

    def __init__(self, path, mode="rb"):
        if path not in AlignmentFile._cache:
            AlignmentFile._cache[path] = self._load(path)
        self.references, self.lengths, self._records = AlignmentFile._cache[path]

    @staticmethod
    def _load(path):
        with gzip.open(path, "rb") as fh:      # BGZF is a series of gzip members
            data = fh.read()
        assert data[:4] == b"BAM\1"
        l_text = struct.unpack_from("<i", data, 4)[0]
        off = 8 + l_text
        n_ref = struct.unpack_from("<i", data, off)[0]
        off += 4
        names, lens = [], []
        for _ in range(n_ref):
            l_name = struct.unpack_from("<i", data, off)[0]
            names.append(data[off + 4:off + 4 + l_name - 1].decode("ascii"))
            lens.append(struct.unpack_from("<i", data, off + 4 + l_name)[0])
            off += 8 + l_name
        recs = []
        n = len(data)
        while off < n:
            block_size = struct.unpack_from("<i", data, off)[0]
            (ref_id, pos, l_read_name, mapq, _bin, n_cigar, flag, l_seq,
             _nref, _npos, _tlen) = struct.unpack_from("<iiBBHHHiiii", data, off + 4)
            p = off + 36
            r = AlignedSegment()
            r.reference_id = ref_id
            r.reference_start = pos
            r.mapping_quality = mapq
            r.flag = flag
            r.query_name = data[p:p + l_read_name - 1].decode("ascii")
            p += l_read_name
            if n_cigar:
                raw = struct.unpack_from("<%dI" % n_cigar, data, p)
                r.cigartuples = [(c & 0xF, c >> 4) for c in raw]
            else:
                r.cigartuples = None
            p += 4 * n_cigar
            nb = (l_seq + 1) >> 1
            packed = data[p:p + nb]
            s = []
            for b in packed:
                s.append(_NT16[b >> 4]); s.append(_NT16[b & 0xF])
            r._seq = "".join(s[:l_seq]) if l_seq else None
            p += nb
            q = data[p:p + l_seq]
            r._qual = None if (l_seq == 0 or q[0] == 0xFF) else bytes(q)
            p += l_seq
            r._aux = _parse_aux(data[p:off + 4 + block_size])
            # bam_endpos (sam.c:347-353)
            if not (flag & 4) and n_cigar > 0:
                rlen = sum(ln for op, ln in r.cigartuples if _CONSUMES_REF[op])
                r.reference_end_htslib = pos + rlen
            else:
                r.reference_end_htslib = pos + 1
            recs.append(r)
            off += 4 + block_size
        return names, lens, recs

    def fetch(self, contig=None, start=None, end=None, until_eof=False):
        if contig is None:
            return iter(self._records)
        tid = self.references.index(contig)
        beg = 0 if start is None else start
        stop = (1 << 29) if end is None else end

        def gen():
            for r in self._records:
                if r.reference_id == tid and r.reference_start < stop and r.reference_end_htslib > beg:
                    yield r
        return gen()

    def count(self, contig=None, start=None, end=None):
        return sum(1 for _ in self.fetch(contig, start, end))

    def close(self):
        pass


# ----------------------------------------------------------------------------------------------------
# VariantFile: what /root/reference/src/utils/pileupGenerator.py:40-62,104-112 touches (fetch(region=), str(rec),
# rec.qual / alleles / ref / start / stop / pos / chrom).  Semantics from the vendored htslib the real pysam wraps:
# QUAL "." -> missing (dep/htslib/vcf.c: bcf_float_set_missing -> pysam None); ALT "." -> no alternate allele;
# rlen = INFO/END - POS + 1 when END is present, else len(REF) (dep/htslib/vcf.c, vcf_parse: "END" handling), and
# pysam's rec.start = POS - 1, rec.stop = start + rlen.
class VariantRecord(object):
    def __init__(self, line):
        self._line = line.rstrip("\n")
        f = self._line.split("\t")
        self.chrom = f[0]
        self.pos = int(f[1])
        self.start = self.pos - 1
        self.id = None if f[2] == "." else f[2]
        self.ref = f[3]
        alts = () if f[4] == "." else tuple(f[4].split(","))
        self.alts = alts or None
        self.alleles = (self.ref,) + alts
        self.qual = None if f[5] == "." else float(f[5])
        rlen = len(self.ref)
        for kv in f[7].split(";"):
            if kv.startswith("END="):
                rlen = int(kv[4:]) - self.pos + 1
        self.rlen = rlen
        self.stop = self.start + rlen

    def __str__(self):
        return self._line + "\n"


class VariantFile(object):
    def __init__(self, path, mode="r"):
        op = gzip.open if str(path).endswith(".gz") else open
        with op(path, "rt") as fh:
            self._records = [VariantRecord(ln) for ln in fh if ln.strip() and not ln.startswith("#")]

    def fetch(self, contig=None, start=None, stop=None, region=None):
        name = region if region is not None else contig
        lo, hi = start, stop
        if name is not None and ":" in name:
            name, span = name.split(":", 1)
            a, b = span.replace(",", "").split("-")
            lo, hi = int(a) - 1, int(b)
        for r in self._records:
            if name is not None and r.chrom != name:
                continue
            if lo is not None and r.stop <= lo:
                continue
            if hi is not None and r.start >= hi:
                continue
            yield r
