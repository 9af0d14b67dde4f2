"""TEST INFRASTRUCTURE ONLY -- runs the UNMODIFIED reference encoder / model from /root/reference.

Only usable where /root/reference is mounted (this build container).  Used by
`oracle/make_golden.py` to produce the committed fixtures under tests/golden/ and by the
`not gpu` tests that pin the restatement (oracle/pileup_oracle.py, oracle/pileup_oracle.c) against
the reference itself.  Never imported by the product package.
"""
import os
import re
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_ROOT = os.environ.get("DEEPORE_REFERENCE", "/root/reference")
REF_BAM = os.path.join(REF_ROOT, "data", "aln_NA12878.pg.chr3_100k-200k.bam")
REF_VCF = os.path.join(REF_ROOT, "data", "vcf_NA12878_S1.small.vcf")
FIXTURE_DIR = os.path.join(HERE, "_ref")
FIXTURE_FASTA = os.path.join(FIXTURE_DIR, "ref_hg19.chr3.9mb.reconstructed.fa")


def reference_available():
    return os.path.isfile(os.path.join(REF_ROOT, "src", "utils", "Pileup.py"))


def _install_standins():
    sd = os.path.join(HERE, "standins")
    if sd not in sys.path:
        sys.path.insert(0, sd)


def import_reference_pileup():
    """Import /root/reference/src/utils/Pileup.py unmodified (module name `Pileup`)."""
    _install_standins()
    p = os.path.join(REF_ROOT, "src", "utils")
    if p not in sys.path:
        sys.path.insert(0, p)
    import Pileup  # noqa
    return Pileup


def import_reference_model():
    """Import /root/reference/src/modules/model.py unmodified (module name `modules.model`)."""
    p = os.path.join(REF_ROOT, "src")
    if p not in sys.path:
        sys.path.insert(0, p)
    import importlib
    return importlib.import_module("modules.model")


_MD_RE = re.compile(r"(\d+)|(\^[A-Za-z]+)|([A-Za-z])")


def reconstruct_reference(bam_path=REF_BAM, lo=99_000, hi=201_000):
    """Rebuild the missing FASTA blob (.MISSING_LARGE_BLOBS) for chr3:[lo,hi) from MD:Z tags.

    Returns a bytearray of length hi-lo filled with 'N' where no read with an MD tag covers."""
    _install_standins()
    import pysam
    ref = bytearray(b"N" * (hi - lo))
    sam = pysam.AlignmentFile(bam_path, "rb")
    for r in sam.fetch():
        if not r.cigartuples or not r.has_tag("MD") or r._seq is None:
            continue
        md = [m for m in _MD_RE.finditer(r.get_tag("MD"))]
        # expand MD into a per-reference-base list over M/=/X and D positions
        mdseq = []            # entries: None (= read base) or a reference letter
        for m in md:
            if m.group(1) is not None:
                mdseq.extend([None] * int(m.group(1)))
            elif m.group(2) is not None:
                mdseq.extend(list(m.group(2)[1:].upper()))
            else:
                mdseq.append(m.group(3).upper())
        pos = r.reference_start
        qi = 0
        k = 0
        seq = r._seq
        ok = True
        for op, n in r.cigartuples:
            if op in (0, 7, 8):
                for i in range(n):
                    if k >= len(mdseq):
                        ok = False
                        break
                    c = mdseq[k] if mdseq[k] is not None else seq[qi]
                    if lo <= pos < hi:
                        ref[pos - lo] = ord(c)
                    pos += 1; qi += 1; k += 1
            elif op == 2:
                for i in range(n):
                    if k >= len(mdseq):
                        ok = False
                        break
                    if lo <= pos < hi:
                        ref[pos - lo] = ord(mdseq[k])
                    pos += 1; k += 1
            elif op == 3:
                pos += n
            elif op in (1, 4):
                qi += n
            if not ok:
                break
    return ref


def ensure_fixture_fasta(path=FIXTURE_FASTA, total=9_000_000, lo=99_000, hi=201_000):
    """Write the reconstructed chr3 FASTA (N-padded to 9 Mb, line width 60 like
    data/ref_hg19.chr3.9mb.fa.fai:1) under oracle/_ref/ (git-ignored)."""
    if os.path.isfile(path):
        return path
    os.makedirs(os.path.dirname(path), exist_ok=True)
    seg = reconstruct_reference(lo=lo, hi=hi)
    full = bytearray(b"N" * total)
    full[lo:hi] = seg
    with open(path + ".tmp", "w") as fh:
        fh.write(">chr3\n")
        s = full.decode("ascii")
        for i in range(0, total, 60):
            fh.write(s[i:i + 60]); fh.write("\n")
    os.replace(path + ".tmp", path)
    return path


def run_reference_window(Pileup, bam, fasta, chromosome, position, flank=100, label=None,
                         variantLengths=None, tmpdir=None, **kw):
    """One unmodified `PileUpGenerator.generatePileup` call (Pileup.py:418-451).

    Returns (uint8[H,W,4] image as stored in the PNG, output label str)."""
    from PIL import Image
    own = tmpdir is None
    if own:
        tmpdir = tempfile.mkdtemp(prefix="dpref_")
    out = os.path.join(tmpdir, "w%s_%d" % (chromosome, position))
    gen = Pileup.PileUpGenerator(bam, fasta)
    lab = gen.generatePileup(chromosome=chromosome, position=position, flankLength=flank,
                             outputFilename=out, label=label if label is not None else "1" * (2 * flank + 1),
                             variantLengths=variantLengths if variantLengths is not None else {}, **kw)
    img = np.array(Image.open(out + ".png"))          # (H, W, 4) : [row=column index][depth][RGBA]
    os.remove(out + ".png")
    if own:
        os.rmdir(tmpdir)
    return img, lab


def qualifying_vcf_positions(vcf_path=REF_VCF, quality_cutoff=60):
    """The window list of pileupGenerator.py:104-105 (QUAL>60, genotype class != 1), parsed as text
    (pysam.VariantFile is absent).  Returns list of 1-based rec.pos."""
    out = []
    with open(vcf_path) as fh:
        for line in fh:
            if line.startswith("#"):
                continue
            f = line.rstrip("\n").split("\t")
            if f[5] == ".":
                continue
            if float(f[5]) <= quality_cutoff:
                continue
            gt = f[-1].split(":")[0].replace("/", "|").split("|")
            if gt[0] == ".":
                continue
            if gt[0] == gt[-1] and gt[0] == "0":
                continue
            out.append(int(f[1]))
    return out
