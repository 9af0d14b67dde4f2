"""dp_bam_scan_static -- the record scan with its scratch in a file-scope device array, for callers that must not
allocate (include/deepore_b200.h; no counterpart in the reference) -- against dp_bam_scan on the same device buffers,
through the C-ABI, on two calls in a row whose inputs differ (the second call sees the first call's scratch contents).
Integer work: the outputs must be identical.  Collected last on purpose (the file name)."""
import ctypes as C

import numpy as np
import pytest

import golden_util as gu

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


def _scan(fn, d_bam, contig, max_records, dev):
    from deepore_b200 import _lib
    n_bytes = int(d_bam.numel())
    name = torch.tensor(list(contig), dtype=torch.uint8, device=dev) if contig else torch.zeros(1, dtype=torch.uint8, device=dev)
    outs = [torch.full((max_records + 1,), -7, dtype=torch.int64, device=dev) for _ in range(4)]
    counts = torch.full((8,), -7, dtype=torch.int64, device=dev)
    st = torch.cuda.current_stream(dev)
    _lib.check(fn(d_bam.data_ptr(), n_bytes, name.data_ptr(), len(contig), max_records, outs[0].data_ptr(), outs[1].data_ptr(),
                  outs[2].data_ptr(), outs[3].data_ptr(), counts.data_ptr(), C.c_void_p(st.cuda_stream)), "scan")
    torch.cuda.synchronize(dev)
    return [o.cpu().numpy() for o in outs], counts.cpu().numpy()


def _streams():
    from deepore_b200 import synth
    ref = synth.random_reference(150_000, seed=31)
    rb = synth.simulate_reads(ref, coverage=30, seed=32)
    big = gu.bam_bytes(rb)                                  # several MB: tens of 64 KB segments
    small = gu.bam_bytes(gu.load("nanopore_small").reads)   # one segment
    return (big, rb.n_reads), (small, None)


def test_static_scan_matches_allocating_scan_twice_in_a_row():
    from deepore_b200 import _lib
    L = _lib.lib()
    dev = torch.device("cuda:0")
    (big, n_big), (small, _) = _streams()
    assert len(big) > 4 * 65536
    for data, contig in ((big, b"chr3"), (small, b"chr3"), (big, b""), (small, b"nope")):
        d_bam = torch.from_numpy(np.frombuffer(data, dtype=np.uint8).copy()).to(dev)
        cap = len(data) // 36 + 1
        want, want_counts = _scan(L.dp_bam_scan, d_bam, contig, cap, dev)
        got, got_counts = _scan(L.dp_bam_scan_static, d_bam, contig, cap, dev)
        assert np.array_equal(got_counts[:7], want_counts[:7]), (contig, got_counts, want_counts)
        kept = int(want_counts[0])
        for g, w in zip(got, want):
            assert np.array_equal(g[:kept + 1], w[:kept + 1])
        if contig == b"chr3":
            assert want_counts[6] == 0 and kept > 0
            if data is big:
                assert kept == n_big
        if contig == b"nope":
            assert want_counts[6] == -3


def test_static_scan_rejects_streams_beyond_its_scratch():
    from deepore_b200 import _lib
    L = _lib.lib()
    dev = torch.device("cuda:0")
    buf = torch.zeros(64, dtype=torch.int64, device=dev)
    st = torch.cuda.current_stream(dev)
    too_big = (1 << 30) + 1                                 # DP_BAM_SCAN_STATIC_MAX_BYTES + 1: refused before any launch
    rc = L.dp_bam_scan_static(buf.data_ptr(), too_big, buf.data_ptr(), 0, 1, buf.data_ptr(), buf.data_ptr(), buf.data_ptr(),
                              buf.data_ptr(), buf.data_ptr(), C.c_void_p(st.cuda_stream))
    assert rc != 0
    with pytest.raises(Exception, match="static scratch"):
        _lib.check(rc, "scan")
    torch.cuda.synchronize(dev)
