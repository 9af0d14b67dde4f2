"""Pileup encoder -- the reference's Python API over the CUDA kernels.

Mirrors /root/reference/src/utils/Pileup.py:
  * `PileUpGenerator(alignmentFile, referenceFile).generatePileup(...)`  (Pileup.py:413-451): same
    signature, defaults, side effect (writes <outputFilename>.png) and return value (the insert-expanded
    label string); the dead `decodeRGB` round trip of :447 is not executed (its result is discarded by the
    reference too).
  * `encode_windows(...)`: the batched form the B200 path actually wants -- all windows of an interval in
    one launch, tensors stay on the device.
Label editing (labelInsertRegion/getOutputLabel, Pileup.py:320-333,374-378) is host string work driven
by the row map the kernel returns.
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, Optional, Tuple

import numpy as np

from . import _lib
from .reads import ReadBatch, BamFile, FastaFile


def _reads_struct(dev: dict, rb: ReadBatch) -> _lib.Reads:
    return _lib.Reads(rb.n_reads, rb.n_cigar, dev["pos"].data_ptr(), dev["flag"].data_ptr(), dev["mapq"].data_ptr(),
                      dev["l_seq"].data_ptr(), dev["cigar_off"].data_ptr(), dev["cigar"].data_ptr(),
                      dev["seq_off"].data_ptr(), dev["seq4"].data_ptr(), dev["qual_off"].data_ptr(),
                      dev["qual"].data_ptr())


class EncoderPlan:
    """Device-resident inputs of one contig interval + reusable workspace; `run()` launches the kernels."""

    def __init__(self, reads: ReadBatch, ref: np.ndarray, ref_base: int, flank: int = 100, coverageCutoff: int = 200,
                 windowCutoff: int = 150, mapQualityCutoff: int = 20, device="cuda"):
        import torch
        _lib.require_cuda()
        self.torch = torch
        self.device = torch.device(device)
        self.reads = reads
        self.params = _lib.EncodeParams(flank, coverageCutoff, windowCutoff, mapQualityCutoff)
        self.flank, self.W, self.H = flank, coverageCutoff, windowCutoff
        self.ref_base = int(ref_base)
        self.ref_len = int(ref.shape[0])
        with torch.cuda.device(self.device):
            self.dev = reads.to_device(self.device)
            self.ref_dev = torch.from_numpy(np.ascontiguousarray(ref, dtype=np.uint8)).pin_memory().to(self.device, non_blocking=True)
            self.rstruct = _reads_struct(self.dev, reads)
            nbytes = _lib.lib().dp_encode_workspace_bytes(reads.n_reads, reads.n_cigar, 0, C.byref(self.params))
            self.workspace = torch.empty(max(int(nbytes), 256), dtype=torch.uint8, device=self.device)

    def run(self, win_qs, win_lo, win_hi, out_img=None, want_rowmap=True, stream=None):
        """win_* : int32 device tensors [Nw].  Returns (img uint8[Nw,H,W,4], rowpos, rowins, status)."""
        torch = self.torch
        nw = int(win_qs.shape[0])
        if out_img is None:
            out_img = torch.empty((nw, self.H, self.W, 4), dtype=torch.uint8, device=self.device)
        rowpos = torch.empty((nw, self.H), dtype=torch.int32, device=self.device) if want_rowmap else None
        rowins = torch.empty((nw, self.H), dtype=torch.int32, device=self.device) if want_rowmap else None
        status = torch.empty((nw,), dtype=torch.int32, device=self.device)
        st = stream if stream is not None else torch.cuda.current_stream(self.device)
        rc = _lib.lib().dp_encode_windows(
            C.byref(self.rstruct), nw, win_qs.data_ptr(), win_lo.data_ptr(), win_hi.data_ptr(),
            self.ref_dev.data_ptr(), self.ref_base, self.ref_len, C.byref(self.params),
            out_img.data_ptr(), rowpos.data_ptr() if want_rowmap else None,
            rowins.data_ptr() if want_rowmap else None, status.data_ptr(),
            self.workspace.data_ptr(), self.workspace.numel(), C.c_void_p(st.cuda_stream))
        _lib.check(rc, "dp_encode_windows")
        return out_img, rowpos, rowins, status


def encode_windows(reads: ReadBatch, win_qs: np.ndarray, ref: np.ndarray, ref_base: int, flank: int = 100,
                   coverageCutoff: int = 200, windowCutoff: int = 150, mapQualityCutoff: int = 20, device="cuda",
                   want_rowmap: bool = True):
    """Host-buffer convenience: uploads, encodes every window, returns device tensors
    (img uint8[Nw,H,W,4], rowpos int32[Nw,H], rowins int32[Nw,H], status int32[Nw])."""
    import torch
    plan = EncoderPlan(reads, ref, ref_base, flank, coverageCutoff, windowCutoff, mapQualityCutoff, device)
    win_qs = np.ascontiguousarray(win_qs, dtype=np.int32)
    lo, hi = reads.window_ranges(win_qs, flank)
    dev = plan.device
    q = torch.from_numpy(win_qs).to(dev)
    l = torch.from_numpy(lo).to(dev)
    h = torch.from_numpy(hi).to(dev)
    return plan.run(q, l, h, want_rowmap=want_rowmap)


_STATUS_EXC = {_lib.DP_EKEY: KeyError, _lib.DP_EINDEX: IndexError, _lib.DP_EINVAL: ValueError}


def expand_label(label: str, variantLengths: Dict[int, int], rowins_full: np.ndarray, windowCutoff: int,
                 noneLabel: str = "0") -> str:
    """labelInsertRegion (Pileup.py:320-333) + getOutputLabel (:374-378), driven by the insert-row layout.

    rowins_full: for every image row IN ORDER (not truncated to windowCutoff) 0 for an anchor row or k+1."""
    c = 0
    i = 0
    n_rows = len(rowins_full)
    while i < n_rows:
        # anchor row i at running index c; count its insert rows
        n = 0
        j = i + 1
        while j < n_rows and rowins_full[j] != 0:
            n += 1
            j += 1
        if n:
            cc = c + 1
            if cc in variantLengths:
                l = variantLengths[cc] - 1
                if l >= n:
                    ins = label[cc - 1] * n
                else:
                    ins = label[cc - 1] * l + noneLabel * (n - l)
            else:
                ins = noneLabel * n
            label = label[:cc] + ins + label[cc:]
            c += n
        c += 1
        i = j
    blank = windowCutoff - len(label)
    return label + noneLabel * blank


class PileUpGenerator:
    """Drop-in for Pileup.PileUpGenerator (Pileup.py:408-451)."""

    _bam_cache: Dict[str, BamFile] = {}
    _fa_cache: Dict[str, FastaFile] = {}
    _batch_cache: Dict[Tuple[str, str], ReadBatch] = {}

    def __init__(self, alignmentFile, referenceFile):
        if alignmentFile not in PileUpGenerator._bam_cache:
            PileUpGenerator._bam_cache[alignmentFile] = BamFile(alignmentFile)
        if referenceFile not in PileUpGenerator._fa_cache:
            PileUpGenerator._fa_cache[referenceFile] = FastaFile(referenceFile)
        self.alignmentFile = alignmentFile
        self.sam = PileUpGenerator._bam_cache[alignmentFile]
        self.fasta = PileUpGenerator._fa_cache[referenceFile]

    def _batch(self, contig: str) -> ReadBatch:
        key = (self.alignmentFile, contig)
        if key not in PileUpGenerator._batch_cache:
            PileUpGenerator._batch_cache[key] = self.sam.read_contig(contig)
        return PileUpGenerator._batch_cache[key]

    def generatePileupTensor(self, chromosome, position, flankLength, coverageCutoff=200, mapQualityCutoff=20,
                             windowCutoff=150):
        """One window -> (uint8[H,W,4] numpy image, rowpos, rowins).  Rows past windowCutoff are reported in
        rowins_full via a second, taller launch only when the label needs them."""
        contig = "chr" + str(chromosome)
        rb = self._batch(contig)
        ref = self.fasta.contig(contig)
        qs = position - flankLength
        L = 2 * flankLength + 1
        lo_g = max(0, qs)
        hi_g = min(ref.shape[0], qs + L + 1)
        img, rowpos, rowins, status = encode_windows(rb, np.asarray([qs], np.int32), ref[lo_g:hi_g], lo_g, flankLength,
                                                     coverageCutoff, windowCutoff, mapQualityCutoff)
        st = int(status[0].item())
        if st != 0:
            raise _STATUS_EXC.get(st, RuntimeError)("window at %s:%d raised status %d" % (contig, position, st))
        return img[0].cpu().numpy(), rowpos[0].cpu().numpy(), rowins[0].cpu().numpy()

    def generatePileup(self, chromosome, position, flankLength, outputFilename, label, variantLengths,
                       forceCoverage=False, coverageCutoff=200, mapQualityCutoff=20, windowCutoff=150,
                       sortColumns=False):
        from PIL import Image
        chromosome = str(chromosome)
        img, rowpos, rowins = self.generatePileupTensor(chromosome, position, flankLength, coverageCutoff,
                                                        mapQualityCutoff, windowCutoff)
        fn = outputFilename if outputFilename.endswith(".png") else outputFilename + ".png"
        Image.fromarray(img, "RGBA").save(fn, "PNG")
        # the label needs the insert layout of ALL rows, also those past windowCutoff: encode with the tallest
        # image the kernel supports when the window is cut
        L = 2 * flankLength + 1
        if (rowpos >= 0).sum() == windowCutoff:
            _, rp2, ri2 = self.generatePileupTensor(chromosome, position, flankLength, 4, mapQualityCutoff, 256)
            nrows = int((rp2 >= 0).sum())
            rowins_full = ri2[:nrows]
        else:
            nrows = int((rowpos >= 0).sum())
            rowins_full = rowins[:nrows]
        return expand_label(label, variantLengths, rowins_full, windowCutoff)
