"""TEST INFRASTRUCTURE ONLY -- ctypes front end of oracle/pileup_oracle.c (the CPU restatement of
/root/reference/src/utils/Pileup.py).  Importable only from tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs.  Never imported by the product package.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "_ref", "libpileup_oracle.so")

# Pileup.calculateQuality (Pileup.py:166-171), evaluated with the reference's own expression
QUALITY_LUT = np.array([int(round((1 - (10 ** ((q) / -10))) * 255)) for q in range(256)], dtype=np.uint8)


class _Reads(C.Structure):
    _fields_ = [("n_reads", C.c_int32), ("n_cigar", C.c_int64), ("pos", C.c_void_p), ("flag", C.c_void_p), ("mapq", C.c_void_p),
                ("l_seq", C.c_void_p), ("cigar_off", C.c_void_p), ("cigar", C.c_void_p),
                ("seq_off", C.c_void_p), ("seq4", C.c_void_p), ("qual_off", C.c_void_p), ("qual", C.c_void_p),
                ("n_bases", C.c_int64)]


class _Params(C.Structure):
    _fields_ = [("flank", C.c_int32), ("coverage_cutoff", C.c_int32), ("window_cutoff", C.c_int32),
                ("map_quality_cutoff", C.c_int32)]


def build(force=False):
    deps = [os.path.join(HERE, "pileup_oracle.c"), os.path.join(HERE, "..", "include", "deepore_b200.h")]
    if force or not os.path.isfile(LIB) or any(os.path.getmtime(LIB) < os.path.getmtime(d) for d in deps):
        subprocess.check_call(["make", "-C", HERE, "-s"])
    return LIB


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(LIB)
        _lib.dpo_encode_windows.restype = C.c_int
    return _lib


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


def encode_windows(reads, win_qs, win_lo, win_hi, ref, ref_base, flank=100, coverage_cutoff=200,
                   window_cutoff=150, map_quality_cutoff=20, want_rowmap=False):
    """reads: object with numpy attributes pos, flag, mapq, l_seq, cigar_off, cigar, seq_off, seq4,
    qual_off, qual.  Returns (img uint8[Nw,H,W,4], status int32[Nw]) (+ rowpos,rowins if asked)."""
    keep = []

    def c(a, dt):
        a = np.ascontiguousarray(a, dtype=dt)
        if a.size == 0:
            a = np.zeros(1, dtype=dt)
        keep.append(a)
        return _ptr(a)

    R = _Reads(int(reads.pos.shape[0]), int(reads.cigar.shape[0]), c(reads.pos, np.int32), c(reads.flag, np.uint16), c(reads.mapq, np.uint8),
               c(reads.l_seq, np.int32), c(reads.cigar_off, np.int64), c(reads.cigar, np.uint32),
               c(reads.seq_off, np.int64), c(reads.seq4, np.uint8), c(reads.qual_off, np.int64),
               c(reads.qual, np.uint8), int(reads.qual.shape[0]))
    P = _Params(flank, coverage_cutoff, window_cutoff, map_quality_cutoff)
    win_qs = np.ascontiguousarray(win_qs, dtype=np.int32)
    win_lo = np.ascontiguousarray(win_lo, dtype=np.int32)
    win_hi = np.ascontiguousarray(win_hi, dtype=np.int32)
    ref = np.ascontiguousarray(ref, dtype=np.uint8)
    nw = int(win_qs.shape[0])
    img = np.empty((nw, window_cutoff, coverage_cutoff, 4), dtype=np.uint8)
    status = np.zeros(nw, dtype=np.int32)
    rowpos = np.empty((nw, window_cutoff), dtype=np.int32) if want_rowmap else None
    rowins = np.empty((nw, window_cutoff), dtype=np.int32) if want_rowmap else None
    lib().dpo_encode_windows(C.byref(R), C.c_int32(nw), _ptr(win_qs), _ptr(win_lo), _ptr(win_hi),
                             _ptr(ref), C.c_int64(ref_base), C.c_int64(ref.shape[0]), C.byref(P),
                             _ptr(QUALITY_LUT), _ptr(img),
                             _ptr(rowpos) if want_rowmap else None, _ptr(rowins) if want_rowmap else None,
                             _ptr(status))
    if want_rowmap:
        return img, status, rowpos, rowins
    return img, status
