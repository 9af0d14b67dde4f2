"""Bounded, seeded slices of the randomised checkers (scripts/enc_stress.py, scripts/model_stress.py) inside the GPU suite, so
that every driver run exercises them: encoder vs the C oracle on randomised workloads (read length 60 bp .. 30 kb, coverage
3 .. 150, indel rates to 12 %, heavy soft clipping, N bases, strides 1 .. 250) and the model forward vs the layer-by-layer
numpy float64 oracle for other weight seeds and random legal pileup tensors."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(script, *args):
    p = subprocess.run([sys.executable, os.path.join(ROOT, "scripts", script)] + [str(a) for a in args], capture_output=True,
                       text=True, timeout=900)
    assert p.returncode == 0, p.stdout[-3000:] + p.stderr[-2000:]
    return p.stdout


def test_encoder_randomised_slice_vs_oracle():
    out = _run("enc_stress.py", 12, 7100)
    assert "12 cases, 0 different" in out


def test_model_randomised_slice_vs_oracle():
    out = _run("model_stress.py", 2, 300)
    assert "2 seeds, 0 over tolerance" in out
