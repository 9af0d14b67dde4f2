"""Pins the CPU restatement (oracle/pileup_oracle.c) against outputs of the UNMODIFIED reference encoder
(/root/reference/src/utils/Pileup.py) committed under tests/golden/ by oracle/make_golden.py."""
import numpy as np
import pytest

import golden_util as gu
import pileup_oracle as po


@pytest.mark.parametrize("name", gu.FIXTURES)
def test_oracle_matches_reference_images(name):
    g = gu.load(name)
    lo, hi = g.reads.window_ranges(g.win_qs, 100)
    img, st = po.encode_windows(g.reads, g.win_qs, lo, hi, g.ref, g.ref_base)
    assert np.array_equal(st, g.status)
    ok = g.status == 0
    assert np.array_equal(img[ok], g.img[ok])


def test_oracle_small_image_parameters():
    g = gu.load("quirks")
    wq = g.z["small_win_qs"]
    lo, hi = g.reads.window_ranges(wq, 40)
    img, st = po.encode_windows(g.reads, wq, lo, hi, g.ref, g.ref_base, flank=40, coverage_cutoff=32,
                                window_cutoff=60, map_quality_cutoff=30)
    assert np.array_equal(st, g.z["small_status"])
    assert np.array_equal(img, g.z["small_img"])


def test_quality_lut_matches_reference_expression():
    # Pileup.calculateQuality (Pileup.py:166-171); first values listed in SURVEY.md A.1
    first = [0, 52, 94, 127, 153, 174, 191, 204, 215, 223, 230, 235, 239, 242, 245, 247, 249, 250, 251, 252, 252,
             253, 253, 254, 254, 254, 254, 254, 255]
    assert po.QUALITY_LUT[:29].tolist() == first
    assert (po.QUALITY_LUT[28:] == 255).all()


def test_oracle_loose_ranges_are_harmless():
    g = gu.load("fuzz")
    n = g.reads.n_reads
    lo = np.zeros(len(g.win_qs), np.int32)
    hi = np.full(len(g.win_qs), n, np.int32)
    img, st = po.encode_windows(g.reads, g.win_qs, lo, hi, g.ref, g.ref_base)
    assert np.array_equal(img, g.img)
