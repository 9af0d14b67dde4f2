"""Device-side BAM ingest, CPU half: the DEFLATE core that bgzf_inflate_kernel runs per thread, compiled
for the host by oracle/Makefile (oracle/_ref/libinflate_host.so), against zlib -- the library the reference's htslib
calls in dep/htslib/bgzf.c:456-499 -- and the host block indexer dp_bgzf_index of the C-ABI library."""
import ctypes
import os
import zlib

import numpy as np
import pytest

import golden_util as gu

HERE = os.path.dirname(os.path.abspath(__file__))


def _host_inflate():
    so = os.path.join(HERE, "..", "oracle", "_ref", "libinflate_host.so")
    import subprocess
    subprocess.check_call(["make", "-C", os.path.join(HERE, "..", "oracle"), "-s"])      # mtime-aware: rebuilds after a core edit
    L = ctypes.CDLL(so)
    L.dpo_inflate.restype = ctypes.c_int
    L.dpo_inflate.argtypes = [ctypes.c_char_p, ctypes.c_uint32, ctypes.c_char_p, ctypes.c_uint32,
                              ctypes.POINTER(ctypes.c_uint32)]

    def run(payload: bytes, cap: int):
        out = ctypes.create_string_buffer(max(cap, 1))
        n = ctypes.c_uint32(0)
        rc = L.dpo_inflate(payload, len(payload), out, cap, ctypes.byref(n))
        return rc, out.raw[:n.value]
    return run


def _payloads():
    rng = np.random.default_rng(11)
    text = (b"ACGTTGCA" * 500 + bytes(rng.integers(65, 70, 3000, dtype=np.uint8))) * 3
    noise = bytes(rng.integers(0, 256, 40000, dtype=np.uint8))
    runs = b"\0" * 65000
    bamlike = gu.bam_bytes(gu.load("quirks").reads)[:60000]
    cases = []
    for name, raw in (("text", text), ("noise", noise), ("runs", runs), ("bam", bamlike), ("one", b"x"), ("empty", b"")):
        for level in (0, 1, 6, 9):
            cases.append((name, level, 0, raw))
        cases.append((name, 6, zlib.Z_FIXED, raw))
        cases.append((name, 6, zlib.Z_HUFFMAN_ONLY, raw))
        cases.append((name, 6, zlib.Z_RLE, raw))
    return cases


def _deflate(raw, level, strategy):
    co = zlib.compressobj(level, zlib.DEFLATED, -15, 9, strategy)
    return co.compress(raw) + co.flush()


def test_inflate_core_matches_zlib():
    run = _host_inflate()
    for name, level, strategy, raw in _payloads():
        payload = _deflate(raw, level, strategy)
        assert zlib.decompress(payload, -15) == raw
        rc, got = run(payload, len(raw))
        assert rc == 0 and got == raw, (name, level, strategy, rc)


def test_inflate_core_multi_block_stream():
    # several DEFLATE blocks in one stream (Z_FULL_FLUSH ends a block and emits an empty stored one)
    run = _host_inflate()
    rng = np.random.default_rng(5)
    co = zlib.compressobj(6, zlib.DEFLATED, -15)
    raw, payload = b"", b""
    for k in range(6):
        piece = bytes(rng.integers(97, 101 + 20 * k, 5000, dtype=np.uint8))
        raw += piece
        payload += co.compress(piece) + co.flush(zlib.Z_FULL_FLUSH if k % 2 else zlib.Z_SYNC_FLUSH)
    payload += co.flush()
    rc, got = run(payload, len(raw))
    assert rc == 0 and got == raw


def test_inflate_core_stored_block_after_huffman_block_at_every_bit_offset():
    # a Huffman block followed by stored ones: the decoder's look-ahead must hand whole bytes back at the byte boundary
    run = _host_inflate()
    rng = np.random.default_rng(23)
    for k in range(200):
        a = bytes(rng.integers(65, 65 + 1 + k % 40, 20 + 7 * k, dtype=np.uint8))
        b = bytes(rng.integers(0, 256, 100 + k, dtype=np.uint8))
        co = zlib.compressobj(6 if k % 3 else 1, zlib.DEFLATED, -15, 9, zlib.Z_FIXED if k % 5 == 0 else 0)
        payload = co.compress(a) + co.flush(zlib.Z_SYNC_FLUSH) + _deflate(b, 0, 0)
        assert zlib.decompress(payload, -15) == a + b
        rc, got = run(payload, len(a) + len(b))
        assert rc == 0 and got == a + b, k


def test_inflate_core_rejects_malformed_input_in_bounds():
    run = _host_inflate()
    raw = b"the quick brown fox jumps over the lazy dog " * 200
    payload = _deflate(raw, 6, 0)
    assert run(payload, len(raw) - 1)[0] == -2                     # output capacity (ISIZE) too small
    assert run(payload[:len(payload) // 2], len(raw))[0] == -1     # truncated input
    assert run(b"\x07", 16)[0] == -3                               # BTYPE 3
    assert run(b"\x01\x05\x00\x00\x00hello", 16)[0] == -4          # stored: LEN / NLEN mismatch
    assert run(b"\x03\x00", 16) == (0, b"")                        # empty fixed block
    # a match that reaches before the start of the output: fixed block, length code 257, distance code 0 (distance 1)
    assert run(bytes([0x03, 0x02]), 16)[0] == -7
    rng = np.random.default_rng(3)
    for _ in range(300):                                           # corrupted payloads: any code, never a crash
        p = bytearray(payload)
        for _k in range(3):
            p[int(rng.integers(0, len(p)))] ^= 1 << int(rng.integers(0, 8))
        rc, got = run(bytes(p), len(raw))
        assert rc <= 0 and len(got) <= len(raw)


def test_bgzf_index_host():
    from deepore_b200 import _lib
    L = _lib.lib()
    data = gu.bam_bytes(gu.load("nanopore_small").reads)
    img = gu.bgzf_compress(data, level=6)
    buf = np.frombuffer(img, dtype=np.uint8)
    n = L.dp_bgzf_index(buf.ctypes.data, buf.size, 0, None, None, None, None, None)
    assert n == (len(data) + 0xff00 - 1) // 0xff00 + 1            # data blocks + the EOF block
    in_off = np.zeros(n, np.int64); in_len = np.zeros(n, np.int32)
    out_off = np.zeros(n + 1, np.int64); out_len = np.zeros(n, np.int32); crc = np.zeros(n, np.uint32)
    assert L.dp_bgzf_index(buf.ctypes.data, buf.size, n, in_off.ctypes.data, in_len.ctypes.data, out_off.ctypes.data,
                           out_len.ctypes.data, crc.ctypes.data) == n
    assert out_off[n] == len(data) and out_len[-1] == 0
    for b in range(n):
        raw = zlib.decompress(img[in_off[b]:in_off[b] + in_len[b]], -15)
        assert raw == data[out_off[b]:out_off[b] + out_len[b]] and (zlib.crc32(raw) & 0xffffffff) == crc[b]
    assert L.dp_bgzf_index(buf.ctypes.data, buf.size - 5, 0, None, None, None, None, None) < 0      # truncated file
    bad = buf.copy(); bad[0] = 0
    assert L.dp_bgzf_index(bad.ctypes.data, bad.size, 0, None, None, None, None, None) < 0          # not BGZF
    assert L.dp_bgzf_index(buf.ctypes.data, 0, 0, None, None, None, None, None) == 0                # empty image
