"""ctypes binding of libdeepore_b200.so (the C-ABI in include/deepore_b200.h).

No fallback: if the library cannot be built/loaded, or a compute entry point is called without a CUDA
device, this raises.
"""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libdeepore_b200.so")

DP_OK, DP_EINVAL, DP_ECUDA, DP_EKEY, DP_EINDEX, DP_ENOMEM, DP_EUNSUPPORTED = 0, -1, -2, -3, -4, -5, -6


class DeeporeError(RuntimeError):
    pass


class Reads(C.Structure):
    _fields_ = [("n_reads", C.c_int32), ("n_cigar", C.c_int64), ("pos", C.c_void_p), ("flag", C.c_void_p),
                ("mapq", C.c_void_p), ("l_seq", C.c_void_p), ("cigar_off", C.c_void_p), ("cigar", C.c_void_p),
                ("seq_off", C.c_void_p), ("seq4", C.c_void_p), ("qual_off", C.c_void_p), ("qual", C.c_void_p),
                ("n_bases", C.c_int64)]


class EncodeParams(C.Structure):
    _fields_ = [("flank", C.c_int32), ("coverage_cutoff", C.c_int32), ("window_cutoff", C.c_int32),
                ("map_quality_cutoff", C.c_int32)]


class ModelWeights(C.Structure):
    _fields_ = [("in_ch", C.c_int32), ("width", C.c_int32), ("hidden", C.c_int32), ("n_class", C.c_int32),
                ("seq_len", C.c_int32), ("n_layers", C.c_int32), ("fold_w", C.c_void_p), ("fold_b", C.c_void_p),
                ("gru_w_ih", (C.c_void_p * 2) * 3), ("gru_w_hh", (C.c_void_p * 2) * 3),
                ("gru_b_ih", (C.c_void_p * 2) * 3), ("gru_b_hh", (C.c_void_p * 2) * 3),
                ("fc1_w", C.c_void_p), ("fc1_b", C.c_void_p)]


_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.isfile(LIB_PATH):
        from . import build as _build
        _build.build()
    if not os.path.isfile(LIB_PATH):
        raise DeeporeError("libdeepore_b200.so is missing and could not be built; there is no CPU fallback")
    L = C.CDLL(LIB_PATH)
    L.dp_last_error.restype = C.c_char_p
    L.dp_version.restype = C.c_int
    L.dp_device_count.restype = C.c_int
    L.dp_encode_workspace_bytes.restype = C.c_int64
    L.dp_encode_workspace_bytes.argtypes = [C.c_int32, C.c_int64, C.c_int64, C.c_int32, C.c_void_p]
    L.dp_encode_windows.restype = C.c_int
    L.dp_encode_windows.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64,
                                    C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                    C.c_int64, C.c_void_p]
    L.dp_encode_prepare.restype = C.c_int
    L.dp_encode_prepare.argtypes = [C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_int64, C.c_void_p]
    L.dp_window_ranges.restype = C.c_int
    L.dp_window_ranges.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_int32, C.c_void_p, C.c_int32, C.c_void_p,
                                   C.c_void_p, C.c_void_p, C.c_void_p]
    L.dp_encode_prepared.restype = C.c_int
    L.dp_encode_prepared.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64,
                                     C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                     C.c_int64, C.c_int32, C.c_int32, C.c_void_p]
    L.dp_window_layout.restype = C.c_int
    L.dp_window_layout.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64,
                                   C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]
    L.dp_encode_fallback_count.restype = C.c_int
    L.dp_encode_fallback_count.argtypes = [C.c_void_p, C.c_int32, C.c_int64, C.c_void_p, C.c_void_p]
    L.dp_set_quality_lut.restype = C.c_int
    L.dp_set_quality_lut.argtypes = [C.c_void_p]
    L.dp_model_create.restype = C.c_int
    L.dp_model_create.argtypes = [C.c_void_p, C.c_int64, C.c_int32, C.POINTER(C.c_void_p)]
    L.dp_model_destroy.restype = C.c_int
    L.dp_model_destroy.argtypes = [C.c_void_p]
    L.dp_model_forward_rows.restype = C.c_int
    L.dp_model_forward_rows.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]
    L.dp_model_forward_windows.restype = C.c_int
    L.dp_model_forward_windows.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p]
    L.dp_model_forward_calls.restype = C.c_int
    L.dp_model_forward_calls.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]
    L.dp_vote.restype = C.c_int
    L.dp_vote.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    L.dp_cross_entropy.restype = C.c_int
    L.dp_cross_entropy.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    L.dp_model_peek.restype = C.c_int
    L.dp_model_peek.argtypes = [C.c_void_p, C.c_int32, C.c_int64, C.c_void_p]
    L.dp_model_profile.restype = C.c_int
    L.dp_model_profile.argtypes = [C.c_void_p, C.c_int32]
    L.dp_model_profile_read.restype = C.c_int
    L.dp_model_profile_read.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    L.dp_gemm_bf16.restype = C.c_int
    L.dp_gemm_bf16.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_int64, C.c_int32,
                               C.c_int64, C.c_int64, C.c_int64, C.c_int32, C.c_void_p]
    L.dp_bgzf_index.restype = C.c_int64
    L.dp_bgzf_index.argtypes = [C.c_void_p, C.c_int64, C.c_int64] + [C.c_void_p] * 5
    L.dp_bgzf_inflate.restype = C.c_int
    L.dp_bgzf_inflate.argtypes = [C.c_void_p, C.c_int64] + [C.c_void_p] * 7
    L.dp_bam_scan.restype = C.c_int
    L.dp_bam_scan.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_int32, C.c_int64] + [C.c_void_p] * 6
    L.dp_bam_scan_static.restype = C.c_int
    L.dp_bam_scan_static.argtypes = L.dp_bam_scan.argtypes
    L.dp_bam_extract.restype = C.c_int
    L.dp_bam_extract.argtypes = [C.c_void_p, C.c_int64] + [C.c_void_p] * 12
    L.dp_png_size.restype = C.c_int64
    L.dp_png_size.argtypes = [C.c_int32, C.c_int32]
    L.dp_png_encode.restype = C.c_int
    L.dp_png_encode.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_int64, C.c_void_p]
    _lib = L
    return L


def check(rc: int, what: str = ""):
    if rc == DP_OK:
        return
    msg = lib().dp_last_error().decode("utf-8", "replace")
    raise DeeporeError("%s failed (%d): %s" % (what or "deepore_b200", rc, msg))


def require_cuda():
    import torch
    if not torch.cuda.is_available() or lib().dp_device_count() <= 0:
        raise DeeporeError("deepore_b200 needs a CUDA device (sm_100a); there is no CPU fallback")


def exported_symbols():
    """Names declared in include/deepore_b200.h (used by the not-gpu ABI test)."""
    import re
    hdr = os.path.join(HERE, "..", "include", "deepore_b200.h")
    txt = open(hdr).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(dp_[a-z0-9_]+)\s*\(", txt)))
