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
from .workload import IntervalShard, make_shard


_FIELDS = ("pos", "flag", "mapq", "l_seq", "cigar_off", "cigar", "seq_off", "seq4", "qual_off", "qual")


def _align(x, a=256):
    return (x + a - 1) // a * a


class ShardBuffers:
    """One IntervalShard packed into ONE pinned host buffer and ONE device buffer (256-byte aligned
    sections), so that handing a shard to the GPU is a single cudaMemcpyAsync."""

    def __init__(self, shard: IntervalShard, device="cuda", pin: bool = True):
        import torch
        self.torch = torch
        self.shard = shard
        self.device = torch.device(device)
        rb = shard.reads
        arrays = [(n, getattr(rb, n)) for n in _FIELDS]
        arrays += [("ref", shard.ref), ("win_qs", shard.win_qs), ("win_lo", shard.win_lo), ("win_hi", shard.win_hi)]
        self.off = {}
        total = 0
        for n, a in arrays:
            self.off[n] = total
            total += _align(max(int(a.nbytes), 16))
        self.nbytes = total
        host = torch.empty(total, dtype=torch.uint8)
        if pin and self.device.type == "cuda":
            host = host.pin_memory()
        hv = host.numpy()
        for n, a in arrays:
            b = np.ascontiguousarray(a).view(np.uint8).reshape(-1)
            hv[self.off[n]:self.off[n] + b.shape[0]] = b
        self.host = host
        self.dev = torch.empty(total, dtype=torch.uint8, device=self.device)
        base = self.dev.data_ptr()
        p = {n: base + o for n, o in self.off.items()}
        self.rstruct = _lib.Reads(rb.n_reads, rb.n_cigar, p["pos"], p["flag"], p["mapq"], p["l_seq"], p["cigar_off"],
                                  p["cigar"], p["seq_off"], p["seq4"], p["qual_off"], p["qual"], int(rb.qual.shape[0]))
        self.ptr = p
        self.n_windows = shard.n_windows
        self.ref_base = int(shard.ref_base)
        self.ref_len = int(shard.ref.shape[0])

    def upload(self, non_blocking: bool = True):
        """Host -> device copy of the whole shard on the current stream."""
        self.dev.copy_(self.host, non_blocking=non_blocking)
        return self


class Encoder:
    """Launches the encoder kernels (C-ABI dp_encode_windows) for uploaded shards."""

    def __init__(self, flank: int = 100, coverageCutoff: int = 200, windowCutoff: int = 150,
                 mapQualityCutoff: int = 20, device="cuda"):
        import torch
        _lib.require_cuda()
        self.torch = torch
        self.device = torch.device(device)
        self.params = _lib.EncodeParams(flank, coverageCutoff, windowCutoff, mapQualityCutoff)
        self.flank, self.W, self.H = flank, coverageCutoff, windowCutoff
        self.workspace = None

    def _workspace(self, n_reads, n_cigar, n_bases, n_windows):
        need = int(_lib.lib().dp_encode_workspace_bytes(n_reads, n_cigar, n_bases, n_windows, C.byref(self.params)))
        if self.workspace is None or self.workspace.numel() < need:
            self.workspace = self.torch.empty(max(need, 256), dtype=self.torch.uint8, device=self.device)
        return self.workspace

    def encode(self, sb: ShardBuffers, out_img=None, rowpos=None, rowins=None, status=None, want_rowmap=True,
               stream=None):
        """Returns (img uint8[Nw,H,W,4], rowpos int32[Nw,H], rowins int32[Nw,H], status int32[Nw]), all on
        the device; nothing is synchronised."""
        torch = self.torch
        nw = sb.n_windows
        dev = self.device
        if out_img is None:
            out_img = torch.empty((nw, self.H, self.W, 4), dtype=torch.uint8, device=dev)
        if want_rowmap:
            if rowpos is None:
                rowpos = torch.empty((nw, self.H), dtype=torch.int32, device=dev)
            if rowins is None:
                rowins = torch.empty((nw, self.H), dtype=torch.int32, device=dev)
        if status is None:
            status = torch.empty((nw,), dtype=torch.int32, device=dev)
        ws = self._workspace(sb.rstruct.n_reads, sb.rstruct.n_cigar, sb.rstruct.n_bases, nw)
        st = stream if stream is not None else torch.cuda.current_stream(dev)
        rc = _lib.lib().dp_encode_windows(
            C.byref(sb.rstruct), nw, sb.ptr["win_qs"], sb.ptr["win_lo"], sb.ptr["win_hi"], sb.ptr["ref"],
            sb.ref_base, sb.ref_len, C.byref(self.params), out_img.data_ptr(),
            rowpos.data_ptr() if want_rowmap else None, rowins.data_ptr() if want_rowmap else None,
            status.data_ptr(), ws.data_ptr(), ws.numel(), C.c_void_p(st.cuda_stream))
        _lib.check(rc, "dp_encode_windows")
        self._last = (sb.rstruct.n_reads, sb.rstruct.n_cigar, st)
        return out_img, (rowpos if want_rowmap else None), (rowins if want_rowmap else None), status


    def fallback_count(self) -> int:
        """Windows of the last encode() that the fast kernel handed to the exact general kernel (diagnostic)."""
        n_reads, n_cigar, st = self._last
        out = C.c_int32(0)
        _lib.check(_lib.lib().dp_encode_fallback_count(self.workspace.data_ptr(), n_reads, n_cigar, C.byref(out),
                                                       C.c_void_p(st.cuda_stream)), "dp_encode_fallback_count")
        return int(out.value)


def encode_windows(reads: ReadBatch, win_qs: np.ndarray, ref: np.ndarray, ref_base: int, flank: int = 100,
                   coverageCutoff: int = 200, windowCutoff: int = 150, mapQualityCutoff: int = 20, device="cuda",
                   want_rowmap: bool = True):
    """Host-buffer convenience: packs the shard, uploads, encodes every window, returns device tensors
    (img uint8[Nw,H,W,4], rowpos int32[Nw,H], rowins int32[Nw,H], status int32[Nw])."""
    _lib.require_cuda()
    shard = make_shard(reads, np.ascontiguousarray(ref, dtype=np.uint8), int(ref_base),
                       np.ascontiguousarray(win_qs, dtype=np.int32), flank)
    sb = ShardBuffers(shard, device).upload()
    enc = Encoder(flank, coverageCutoff, windowCutoff, mapQualityCutoff, device)
    return enc.encode(sb, want_rowmap=want_rowmap)


def window_layout(reads: ReadBatch, win_qs: np.ndarray, ref: np.ndarray, ref_base: int, flank: int = 100,
                  coverageCutoff: int = 200, windowCutoff: int = 150, mapQualityCutoff: int = 20, device="cuda"):
    """dp_window_layout: image rows per window column, NOT cut at windowCutoff -> (int32 [Nw, 2*flank+1] on the host: 0 for a
    column without entries, else 1 + its insert rows; status int32 [Nw])."""
    import torch
    _lib.require_cuda()
    shard = make_shard(reads, np.ascontiguousarray(ref, dtype=np.uint8), int(ref_base),
                       np.ascontiguousarray(win_qs, dtype=np.int32), flank)
    sb = ShardBuffers(shard, device).upload()
    enc = Encoder(flank, coverageCutoff, windowCutoff, mapQualityCutoff, device)
    nw = sb.n_windows
    dev = enc.device
    colrows = torch.zeros((max(nw, 1), 2 * flank + 1), dtype=torch.int32, device=dev)
    status = torch.zeros((max(nw, 1),), dtype=torch.int32, device=dev)
    ws = enc._workspace(sb.rstruct.n_reads, sb.rstruct.n_cigar, sb.rstruct.n_bases, nw)
    st = torch.cuda.current_stream(dev)
    with torch.cuda.device(dev):
        rc = _lib.lib().dp_window_layout(C.byref(sb.rstruct), nw, sb.ptr["win_qs"], sb.ptr["win_lo"], sb.ptr["win_hi"],
                                         sb.ptr["ref"], sb.ref_base, sb.ref_len, C.byref(enc.params), colrows.data_ptr(),
                                         status.data_ptr(), ws.data_ptr(), ws.numel(), C.c_void_p(st.cuda_stream))
    _lib.check(rc, "dp_window_layout")
    return colrows[:nw].cpu().numpy(), status[:nw].cpu().numpy()


def rowins_of_layout(colrows_row: np.ndarray) -> np.ndarray:
    """Insert index of every image row of a window IN ORDER (0 for an anchor row, k for its k-th insert row) from the rows per
    column that dp_window_layout returns: what `expand_label` walks."""
    out = []
    for n in colrows_row.tolist():
        if n > 0:
            out.extend(range(n))
    return np.asarray(out, dtype=np.int32)


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


_CATS = "ACGTDMN"                                   # sortingKey order of Pileup.py:66-73 (category index of the kernels)
_RGB_OF = {"M": (255, 255, 255), "A": (255, 0, 0), "C": (255, 255, 0), "G": (0, 255, 0), "T": (0, 0, 255),
           "I": (255, 0, 255), "D": (0, 255, 255), "N": (0, 0, 0)}
_CAT_OF_RGB = {v: k for k, v in _RGB_OF.items()}
_TEXT_OF_BITS = [[[" ", "T"], ["G", "D"]], [["A", "_"], ["C", "."]]]     # Pileup.py:77 (decodeRGB alphabet)


class Pileup:
    """Reference surface of Pileup.Pileup (Pileup.py:21-405) over the CUDA encoder: same constructor arguments, the
    methods generatePileup() calls (`iterateReads`, `generatePileupImage`, `savePileupRGB`, `getOutputLabel`,
    `decodeRGB`) and the attributes `pileupColumns`, `insertColumns`, `pileupImage`, `refSequence`, `label`.

    `sam` / `fasta` are the handles a `PileUpGenerator` of this package holds (`reads.BamFile`, `reads.FastaFile`).
    The window is encoded on the GPU in `iterateReads`; the dictionaries are rebuilt from the kernel's rows for the
    part of the window that reaches the image (the reference keeps every entry, the image keeps the first
    coverageCutoff-1 per column and the first windowCutoff rows -- what the PNG, the label and the model see)."""

    def __init__(self, sam, fasta, chromosome, queryStart, flankLength, outputFilename, label, variantLengths,
                 coverageCutoff, mapQualityCutoff, windowCutoff, sortColumns, subsampleRate=0, forceCoverage=False,
                 arrayInitializationFactor=2):
        self.length = flankLength * 2 + 1
        self.label = label
        self.variantLengths = variantLengths
        self.subsampleRate = subsampleRate
        self.coverageCutoff = coverageCutoff
        self.outputFilename = outputFilename
        self.queryStart = queryStart
        self.queryEnd = queryStart + self.length
        self.chromosome = str(chromosome)
        self.windowCutoff = windowCutoff
        self.mapQualityCutoff = mapQualityCutoff
        self.sortColumns = sortColumns
        self.forceCoverage = forceCoverage
        self.flankLength = flankLength
        self.sam, self.fasta = sam, fasta
        self.noneChar, self.noneLabel = "_", "0"
        self.SNPtoRGB = {k: list(v) for k, v in _RGB_OF.items()}
        self.SNPtoRGB[self.noneChar] = [0, 0, 0]
        contig = "chr" + self.chromosome
        ref = fasta.contig(contig)
        lo, hi = max(0, queryStart), min(ref.shape[0], self.queryEnd + 1)
        self._ref, self._ref_base = ref[lo:hi], lo
        self.refSequence = ref[lo:hi].tobytes().decode("ascii")          # 0-based [queryStart, queryEnd] inclusive (:41)
        self.pileupColumns = dict()
        self.insertColumns = dict()
        self.pileupImage = list()
        self._img = self._rowpos = self._rowins = self._rowins_full = None

    # ---- Pileup.py:152-163 (+ parseRead :174-264): here one CUDA launch for the window
    def iterateReads(self):
        contig = "chr" + self.chromosome
        cache = self.sam.__dict__.setdefault("_dp_batches", {})          # one parsed ReadBatch per contig and handle
        if contig not in cache:
            cache[contig] = self.sam.read_contig(contig)
        rb = cache[contig]
        wq = np.asarray([self.queryStart], np.int32)

        def run(W, H):
            img, rowpos, rowins, status = encode_windows(rb, wq, self._ref, self._ref_base, self.flankLength, W, H,
                                                         self.mapQualityCutoff)
            st = int(status[0].item())
            if st != 0:
                raise _STATUS_EXC.get(st, RuntimeError)("window at chr%s:%d raised status %d"
                                                        % (self.chromosome, self.queryStart + self.flankLength, st))
            return img[0].cpu().numpy(), rowpos[0].cpu().numpy(), rowins[0].cpu().numpy()

        self._img, self._rowpos, self._rowins = run(self.coverageCutoff, self.windowCutoff)
        nrows = int((self._rowpos >= 0).sum())
        self._rowins_full = self._rowins[:nrows]
        if nrows == self.windowCutoff:                                  # rows were cut: the label needs all of them
            colrows, _ = window_layout(rb, wq, self._ref, self._ref_base, self.flankLength, self.coverageCutoff,
                                       self.windowCutoff, self.mapQualityCutoff)
            self._rowins_full = rowins_of_layout(colrows[0])
        # dictionaries of the drawn part: position -> category -> [[R,G,B,q], ...] ; anchor -> [k] -> base -> [...]
        for y in range(nrows):
            pos, k = int(self._rowpos[y]), int(self._rowins[y])
            row = self._img[y]
            filler = (255, 0, 255, 255) if k else (0, 0, 0, 255)
            cells = dict()
            for px in row[1:]:
                t = tuple(int(v) for v in px)
                if t == filler:
                    break
                cells.setdefault(_CAT_OF_RGB[t[:3]], []).append(list(t))
            if k == 0:
                self.pileupColumns[pos] = cells
            else:
                self.insertColumns.setdefault(pos, []).append(cells)

    # ---- Pileup.py:295-317: rows of RGBA tuples, 'I' spliced into refSequence, label widened at insert rows
    def generatePileupImage(self):
        if self._img is None:
            self.iterateReads()
        nrows = int((self._rowpos >= 0).sum())
        self.pileupImage = [[tuple(int(v) for v in px) for px in self._img[y]] for y in range(nrows)]
        c = 0
        i = 0
        full = self._rowins_full
        while i < len(full):
            n = 0
            j = i + 1
            while j < len(full) and full[j] != 0:
                n += 1
                j += 1
            if n:
                self.labelInsertRegion(c + 1, n)
                for _ in range(n):
                    c += 1
                    self.refSequence = self.refSequence[:c] + "I" + self.refSequence[c:]
            c += 1
            i = j

    def labelInsertRegion(self, c, n):                     # Pileup.py:320-333
        if c in self.variantLengths:
            l = self.variantLengths[c] - 1
            ins = self.label[c - 1] * n if l >= n else self.label[c - 1] * l + self.noneLabel * (n - l)
        else:
            ins = self.noneLabel * n
        self.label = self.label[:int(c)] + ins + self.label[int(c):]

    def savePileupRGB(self, filename):                     # Pileup.py:336-361
        from PIL import Image
        if not filename.endswith(".png"):
            filename += ".png"
        self.pileupImage = [row[:self.coverageCutoff] for row in self.pileupImage[:self.windowCutoff]]
        Image.fromarray(np.ascontiguousarray(self._img), "RGBA").save(filename, "PNG")

    def getOutputLabel(self):                              # Pileup.py:374-378
        self.label += self.noneLabel * (self.windowCutoff - len(self.label))
        return self.label

    def decodeRGB(self, filename):                         # Pileup.py:381-405 (debug view; generatePileup discards it)
        from PIL import Image
        px = np.asarray(Image.open(filename).convert("RGBA"))
        bits = (px[..., :3] // 255).astype(np.int64)
        return ["".join(_TEXT_OF_BITS[r][g][b] for r, g, b in row) for row in bits]


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
        """Pileup.py:418-451, same call sequence; the dead decodeRGB round trip of :447 is not executed."""
        queryStart = position - flankLength
        self.sam.__dict__.setdefault("_dp_batches", {})["chr" + str(chromosome)] = self._batch("chr" + str(chromosome))
        pileup = Pileup(self.sam, self.fasta, str(chromosome), queryStart, flankLength, outputFilename, label,
                        variantLengths, windowCutoff=windowCutoff, forceCoverage=forceCoverage,
                        coverageCutoff=coverageCutoff, mapQualityCutoff=mapQualityCutoff, sortColumns=sortColumns)
        pileup.iterateReads()
        pileup.generatePileupImage()
        pileup.savePileupRGB(outputFilename)
        return pileup.getOutputLabel()
