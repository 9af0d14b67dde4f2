"""Window / label producer: API of /root/reference/src/utils/pileupGenerator.py (:23-123) over the CUDA encoder.

The reference walks the VCF, and for every qualifying record (QUAL > 60, genotype class != hom-ref) opens the BAM and the FASTA
again, encodes ONE window, writes its PNG and one line of `summary-<region>.csv` (pileupGenerator.py:97-123).  Here the same
function names, arguments, file names and CSV lines are kept, but the work is laid out for the GPU: the VCF is parsed once on
the host, every window of the run goes to the device in ONE interval shard (`dp_encode_windows`, one H2D copy, one launch group),
the labels are widened from the kernel's insert-row map, and the PNG files are produced on the device (dp_png_encode).

`pysam.VariantFile` is replaced by a plain-text VCF reader (`read_vcf`) with the record semantics the reference relies on:
`qual` None for ".", `alleles` = REF + ALTs ("." = none), `start` = POS-1, `stop` = start + rlen with rlen = INFO/END-POS+1 when
END is present (gVCF blocks) else len(REF) -- dep/htslib/vcf.c, as wrapped by pysam.

No CPU fallback: the windows are encoded by the CUDA library or the call raises.
"""
from __future__ import annotations

import gzip
import os
import sys
from concurrent.futures import ThreadPoolExecutor
from timeit import default_timer as timer
from typing import Dict, List, Optional, Tuple

import numpy as np

# module-level switches of pileupGenerator.py:17-20 (same names, same meaning)
allVariantRecord: Dict[int, Tuple[int, Optional[int], Optional[int]]] = {}
subregion = ''
cutoffOutput = False
cutoff = 350


class VcfRecord:
    """One VCF data line with the attribute names of pysam.VariantRecord that pileupGenerator.py reads."""
    __slots__ = ("chrom", "pos", "start", "stop", "ref", "alleles", "qual", "_line")

    def __init__(self, line: str):
        self._line = line.rstrip("\n")
        f = self._line.split("\t")
        if len(f) < 8:
            raise ValueError("not a VCF data line: %r" % line[:60])
        self.chrom = f[0]
        self.pos = int(f[1])
        self.start = self.pos - 1
        self.ref = f[3]
        self.alleles = (self.ref,) + (() if f[4] == "." else tuple(f[4].split(",")))
        self.qual = None if f[5] == "." else float(f[5])
        rlen = len(self.ref)
        for kv in f[7].split(";"):
            if kv.startswith("END="):
                rlen = int(kv[4:]) - self.pos + 1
        self.stop = self.start + rlen

    def __str__(self):
        return self._line + "\n"


def read_vcf(path: str, region: Optional[str] = None) -> List[VcfRecord]:
    """Records of a plain or gzip/bgzip VCF in file order; `region` = "chr3" or "chr3:100-200" (1-based inclusive, the
    tabix convention of VariantFile.fetch(region=...), pileupGenerator.py:41,104)."""
    name, lo, hi = region, None, None
    if region and ":" in region:
        name, span = region.split(":", 1)
        a, b = span.replace(",", "").split("-")
        lo, hi = int(a) - 1, int(b)
    op = gzip.open if str(path).endswith(".gz") else open
    out = []
    with op(path, "rt") as fh:
        for ln in fh:
            if not ln.strip() or ln.startswith("#"):
                continue
            r = VcfRecord(ln)
            if name and r.chrom != name:
                continue
            if lo is not None and (r.stop <= lo or r.start >= hi):
                continue
            out.append(r)
    return out


def getClassForGenotype(gtField):
    """pileupGenerator.py:23-32: 1 = hom-ref or missing, 2 = heterozygous, 3 = hom-alt."""
    a, b = gtField[0], gtField[-1]
    if a == '.':
        return 1
    if a != b:
        return 2
    return 1 if a == '0' else 3


def getGTField(rec):
    """pileupGenerator.py:35-36: genotype of the LAST sample column, split on / | or backslash."""
    gt = str(rec).rstrip().split('\t')[-1].split(':')[0]
    for sep in ('/', '\\'):
        gt = gt.replace(sep, '|')
    return gt.split('|')


def populateRecordDictionary(vcf_region, vcfFile, qualityCutoff=60):
    """pileupGenerator.py:39-62: reference position -> (genotype class, insert length or None, delete length or None) for
    every position a confident record spans (all classes, hom-ref included)."""
    for rec in read_vcf(vcfFile, "chr" + vcf_region + subregion):
        gtField = getGTField(rec)
        genotypeClass = getClassForGenotype(gtField)
        if rec.qual is None or not rec.qual > qualityCutoff:
            continue
        longest = max([0] + [len(rec.alleles[int(gt)]) for gt in gtField])     # int('.') raises, as in the reference
        entry = (genotypeClass, longest, None) if longest > len(rec.ref) else (genotypeClass, None, longest)
        for i in range(rec.start, rec.stop):
            allVariantRecord[i] = entry
    return allVariantRecord


def getLabel(start, end):
    """pileupGenerator.py:65-95: label draft (one class digit per reference position of [start, end), '1' where no record)
    and the window-relative positions where an insert / a delete of a given allele length begins."""
    digits = []
    insertLengths, deleteLengths = dict(), dict()
    ins_left = del_left = 0
    for j, i in enumerate(range(start, end)):
        rec = allVariantRecord.get(i)
        if rec is None:
            digits.append('1')
            continue
        digits.append(str(rec[0]))
        if ins_left:
            ins_left -= 1
        elif rec[1] is not None:
            insertLengths[j] = int(rec[1])
            ins_left = rec[1] - 1
        if del_left:
            del_left -= 1
        elif rec[2] is not None:
            deleteLengths[j] = int(rec[2])
            del_left = rec[2] - 1
    return ''.join(digits), insertLengths, deleteLengths


def qualifying_records(vcf_region, vcfFile, qualityCutoff=60) -> List[VcfRecord]:
    """The records a window is cut around (pileupGenerator.py:104-105)."""
    return [r for r in read_vcf(vcfFile, "chr" + vcf_region + subregion)
            if r.qual is not None and r.qual > qualityCutoff and getClassForGenotype(getGTField(r)) != 1]


def encode_labelled_windows(bamFile, refFile, chromosome, positions, labels, variantLengths, flankLength=100,
                            coverageCutoff=200, mapQualityCutoff=20, windowCutoff=150, device="cuda", keep_device=False):
    """Batched `PileUpGenerator.generatePileup` (Pileup.py:418-451) without the PNG: windows centred on the 0-based
    `positions` of "chr"+chromosome of a BAM + FASTA pair -> see `encode_labelled_batch`."""
    from .pileup import PileUpGenerator
    gen = PileUpGenerator(bamFile, refFile)
    contig = "chr" + str(chromosome)
    return encode_labelled_batch(gen._batch(contig), gen.fasta.contig(contig), 0, positions, labels, variantLengths,
                                 flankLength, coverageCutoff, mapQualityCutoff, windowCutoff, device, contig, keep_device)


def encode_labelled_batch(rb, ref, ref_base, positions, labels, variantLengths, flankLength=100, coverageCutoff=200,
                          mapQualityCutoff=20, windowCutoff=150, device="cuda", contig="chr?", keep_device=False):
    """Reads of one contig (`reads.ReadBatch`) + reference bytes starting at genomic `ref_base` + N window centres with their
    label drafts and insert-length dictionaries -> (uint8 [N, windowCutoff, coverageCutoff, 4] host array, list of output
    labels, rowpos, rowins).  One shard, one `dp_encode_windows` call; for the windows whose rows were cut at windowCutoff
    a `dp_window_layout` call returns the uncut row layout, because the label is widened by ALL insert rows of the window
    (Pileup.py:309, 320-333).  A window the reference would raise on raises the same exception type here.  `keep_device`: the images stay a
    CUDA tensor (for the device-side PNG encoder, png.write_files) instead of being copied to the host."""
    from . import _lib
    from .pileup import encode_windows, expand_label, _STATUS_EXC
    _lib.require_cuda()
    ref = np.asarray(ref, dtype=np.uint8)
    positions = np.asarray(positions, dtype=np.int64)
    n = int(positions.shape[0])
    L = 2 * flankLength + 1
    if n == 0:
        return (np.zeros((0, windowCutoff, coverageCutoff, 4), np.uint8), [], np.zeros((0, windowCutoff), np.int32),
                np.zeros((0, windowCutoff), np.int32))
    order = np.argsort(positions, kind="stable")                     # the shard wants ascending window starts
    qs = (positions[order] - flankLength).astype(np.int32)
    lo_g = max(int(ref_base), int(qs.min()))
    hi_g = min(int(ref_base) + int(ref.shape[0]), int(qs.max()) + L + 1)
    ref_cut = ref[lo_g - int(ref_base):hi_g - int(ref_base)]

    def run(sel, W, H):
        img, rowpos, rowins, status = encode_windows(rb, qs[sel], ref_cut, lo_g, flankLength, W, H, mapQualityCutoff,
                                                     device=device)
        status = status.cpu().numpy()
        bad = np.nonzero(status != 0)[0]
        if bad.size:
            k = int(sel[bad[0]])
            raise _STATUS_EXC.get(int(status[bad[0]]), RuntimeError)(
                "window at %s:%d raised status %d" % (contig, int(positions[order[k]]), int(status[bad[0]])))
        return (img if keep_device else img.cpu().numpy()), rowpos.cpu().numpy(), rowins.cpu().numpy()

    every = np.arange(n)
    img_s, rowpos_s, rowins_s = run(every, coverageCutoff, windowCutoff)
    nrows = (rowpos_s >= 0).sum(axis=1)
    full = [rowins_s[k, :nrows[k]] for k in range(n)]
    cut = np.nonzero(nrows == windowCutoff)[0]
    if cut.size:                                                     # the label is widened by ALL insert rows of the window
        from .pileup import rowins_of_layout, window_layout
        colrows, _ = window_layout(rb, qs[cut], ref_cut, lo_g, flankLength, coverageCutoff, windowCutoff, mapQualityCutoff,
                                   device=device)
        for a, k in enumerate(cut):
            full[k] = rowins_of_layout(colrows[a])
    inv = np.empty(n, dtype=np.int64)
    inv[order] = every
    out_labels = [expand_label(labels[i], variantLengths[i], full[inv[i]], windowCutoff) for i in range(n)]
    if keep_device:
        import torch
        return img_s[torch.from_numpy(inv).to(img_s.device)], out_labels, rowpos_s[inv], rowins_s[inv]
    return img_s[inv], out_labels, rowpos_s[inv], rowins_s[inv]


def write_pngs(paths, images, threads: Optional[int] = None):
    """RGBA PNG per window, as savePileupRGB stores it (Pileup.py:336-361: 200 px wide, 150 high, pixel (i, j) = row j, depth
    i).  A CUDA tensor is encoded on the device (png.write_files: dp_png_encode, only finished files cross PCIe); host arrays
    -- callers that already hold the pixels on the host -- go through PIL on a thread pool."""
    if hasattr(images, "is_cuda") and images.is_cuda:
        from . import png
        png.write_files(list(paths), images, threads)
        return
    from PIL import Image

    def one(k):
        Image.fromarray(np.ascontiguousarray(images[k]), "RGBA").save(paths[k], "PNG")
    with ThreadPoolExecutor(max_workers=threads or min(32, os.cpu_count() or 1)) as ex:
        list(ex.map(one, range(len(paths))))


def generatePileupBasedonVCF(vcf_region, bamFile, refFile, vcfFile, output_dir, window_size, qualityCutoff=60,
                             device="cuda"):
    """pileupGenerator.py:97-123, same arguments and outputs: `<output_dir><chrom>-<pos>.png` per qualifying record and
    `<output_dir>summary-<vcf_region>.csv` with one `absolute png path,label` line per record, in VCF order.  Returns the
    number of windows written."""
    start_timer = timer()
    populateRecordDictionary(vcf_region, vcfFile)                    # the reference passes no cutoff here either (:100)
    recs = qualifying_records(vcf_region, vcfFile, qualityCutoff)
    if cutoffOutput:
        recs = recs[:cutoff + 1]                                     # the reference breaks once cnt > cutoff (:121-123)
    drafts, inserts = [], []
    for rec in recs:
        lab, ins, _ = getLabel(rec.pos - window_size - 1, rec.pos + window_size)
        drafts.append(lab)
        inserts.append(ins)
    img, labels, _, _ = encode_labelled_windows(bamFile, refFile, vcf_region, [r.pos - 1 for r in recs], drafts, inserts,
                                                flankLength=window_size, device=device, keep_device=True)
    names = [output_dir + r.chrom + "-" + str(r.pos) for r in recs]
    last = {nm: k for k, nm in enumerate(names)}                     # records sharing a position share a file: last one wins
    keep = sorted(last.values())
    if keep:
        import torch
        write_pngs([names[k] + ".png" for k in keep], img[torch.as_tensor(keep, device=img.device)])
    with open(output_dir + 'summary' + '-' + vcf_region + ".csv", 'w') as smry:
        for nm, lab in zip(names, labels):
            smry.write(os.path.abspath(nm) + ".png," + str(lab) + '\n')
    print(str(len(recs)) + " Records done", file=sys.stderr)
    print("TIME elapsed " + str(timer() - start_timer), file=sys.stderr)
    return len(recs)


if __name__ == '__main__':
    import argparse
    parser = argparse.ArgumentParser()                               # flags of pileupGenerator.py:130-179
    parser.add_argument("--bam", type=str, required=True, help="BAM file with alignments.")
    parser.add_argument("--ref", type=str, required=True, help="Reference corresponding to the BAM file.")
    parser.add_argument("--vcf", type=str, required=False, help="VCF file containing SNPs and SVs.")
    parser.add_argument("--region", type=str, default="chr3", help="Site region. Ex: chr3")
    parser.add_argument("--vcf_region", type=str, default="3", help="Site region. Ex: chr3")
    parser.add_argument("--coverage", type=int, default=50, help="Read coverage, default is 50x.")
    parser.add_argument("--output_dir", type=str, default="output/", help="Name of output directory")
    parser.add_argument("--window_size", type=int, default=100, help="Window size of pileup.")
    FLAGS, unparsed = parser.parse_known_args()
    generatePileupBasedonVCF(FLAGS.vcf_region, FLAGS.bam, FLAGS.ref, FLAGS.vcf, FLAGS.output_dir, FLAGS.window_size)
