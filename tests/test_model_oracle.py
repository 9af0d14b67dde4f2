"""CPU tests (no GPU): the numpy model oracle is pinned against outputs of the UNMODIFIED reference model
(tests/golden/model_seed7.npz, written by oracle/make_model_golden.py), and the host-side operator fold is checked
against the oracle's layer-by-layer evaluation."""
import os

import numpy as np
import pytest

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "model_seed7.npz")


@pytest.fixture(scope="module")
def gold():
    return np.load(GOLD)


@pytest.fixture(scope="module")
def state(gold):
    from deepore_b200 import modelio
    return modelio.seeded_state(int(gold["seed"]))


CALLS = [0, 1, 2, 3, 4, 60, 61, 62, 145, 146, 147]


def test_oracle_matches_reference_logits(gold, state):
    import model_oracle as mo
    for wi in (0, 2):
        logits, y = mo.forward_image(state, gold["img"][wi], calls=CALLS, return_y=True)
        ref = gold["logits"][wi][CALLS]
        assert np.abs(logits - ref).max() < 2e-5          # reference is fp32, oracle fp64
        if wi == 0:
            assert np.abs(y - gold["y_window0"][CALLS]).max() < 1e-5


def test_oracle_vote_matches_reference_loop(gold):
    import model_oracle as mo
    for wi in range(gold["pred"].shape[0]):
        assert np.array_equal(mo.argmax_first(gold["logits"][wi]), gold["pred"][wi])
        assert np.array_equal(mo.vote(gold["pred"][wi]), gold["calls"][wi])
    # three-way tie -> smallest class; pairs -> the pair
    p = np.zeros((5, 3), np.uint8)
    p[0, 2], p[1, 1], p[2, 0] = 2, 1, 0          # row 2: votes (2,1,0) -> 0
    p[1, 2], p[2, 1], p[3, 0] = 1, 2, 2          # row 3: votes (1,2,2) -> 2
    v = mo.vote(p)
    assert v[2] == 0 and v[3] == 2 and v[0] == -1 and v[1] == -1 and v[5] == -1 and v[6] == -1


def test_fold_equals_layered_front(gold, state):
    """fold_w/fold_b reproduce cnn_fc_rnn(feat) of the reference for every call of window 0 (fp64 fold, fp32 ref)."""
    from deepore_b200 import fold
    fw, fb = fold.fold_front(state, 4, 200, 3)
    img = gold["img"][0].astype(np.float64)
    calls = np.arange(148)
    x = np.stack([img[r:r + 3] for r in calls])                      # [148,3,200,4]
    y = np.einsum("thswc,nswc->nth", fw.astype(np.float64), x, optimize=True) + fb[None]
    assert np.abs(y - gold["y_window0"]).max() < 1e-5


def test_fold_linearity_and_border_columns(state):
    """The folded operator must equal the layered oracle on inputs that light up single pixels at the borders
    (columns 0, 14, 15, 184, 185, 199), where the zero padding of the intermediate layers matters."""
    import model_oracle as mo
    from deepore_b200 import fold
    fw, fb = fold.fold_front(state, 4, 200, 3)
    rng = np.random.default_rng(3)
    img = np.zeros((3, 200, 4), np.uint8)
    for col in (0, 1, 14, 15, 16, 100, 183, 184, 185, 198, 199):
        img[rng.integers(0, 3), col, rng.integers(0, 4)] = rng.integers(1, 256)
    _, y = mo.forward_image(state, img, calls=[0], return_y=True)
    yf = np.einsum("thswc,swc->th", fw.astype(np.float64), img.astype(np.float64)) + fb
    assert np.abs(yf - y[0]).max() < 1e-6


def test_call_set_golden_is_consistent_with_the_logit_golden(gold):
    """tests/golden/bundled_calls_seed7.npz (all 260 bundled windows through the unmodified reference model) and
    model_seed7.npz (4 of them, with logits) come from separate runs of the reference: same classes, same votes, and
    the stored margins are the logits' top-2 margins."""
    import os
    full = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "bundled_calls_seed7.npz"))
    w = gold["windows"]
    assert np.array_equal(full["pred"][w], gold["pred"])
    assert np.array_equal(full["calls"][w], gold["calls"])
    srt = np.sort(gold["logits"], axis=-1)
    assert np.abs(full["margin"][w] - (srt[..., -1] - srt[..., -2])).max() < 1e-5
    assert full["pred"].shape == (260, 148, 3) and full["calls"].shape == (260, 150)
    assert ((full["calls"][:, 2:148] >= 0).all() and (full["calls"][:, :2] == -1).all()
            and (full["calls"][:, 148:] == -1).all())


def test_torch_restatement_matches_reference_logits(gold, state):
    """oracle/torch_model.py (the torch-CPU leg of bench.py's CPU baseline: nn.Conv2d / BatchNorm2d / Linear / GRU modules
    filled from the numpy state, forward restated from model.py:93-156) against the logits of the unmodified reference
    class -- same torch kernels, so equal up to thread-count dependent summation order."""
    import torch
    import torch_model as tm
    torch.set_num_threads(max(1, min(8, os.cpu_count() or 1)))
    m = tm.build(state)
    got = tm.forward_image_rows(m, gold["img"][:2], CALLS)
    assert got.shape == (2, len(CALLS), 3, 3)
    assert np.abs(got - gold["logits"][:2][:, CALLS]).max() < 2e-5


TRAINED = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "bundled_calls_trained.npz")


@pytest.mark.skipif(not os.path.isfile(TRAINED), reason="tests/golden/bundled_calls_trained.npz not generated")
def test_trained_checkpoint_golden_is_reproduced_by_the_torch_restatement():
    """SURVEY 8(f4) slice: the goldens of the SGD-stepped checkpoint (oracle/make_trained_golden.py, unmodified reference
    model) are reproduced by the restated torch model from seeded_state + the committed small tensors -- i.e. the fixture
    holds everything the training changed -- and the steps did move the model: fewer calls disagree with the labels."""
    import golden_util as gu
    import torch
    import torch_model as tm
    torch.set_num_threads(max(1, min(8, os.cpu_count() or 1)))
    st, z = gu.trained_state(TRAINED)
    changed = [k for k in z.files if k.startswith("state/")]
    assert "state/fc1.weight" in changed and "state/batchNorm.running_mean" in changed and int(z["steps"]) > 0
    assert sum(z[k].size for k in changed) < 20000
    g = gu.load("bundled_na12878")
    rows = [0, 1, 73, 146, 147]
    m = tm.build(st)
    got = tm.forward_image_rows(m, g.img[:2], rows)
    assert np.abs(got - z["logits"][:2][:, rows]).max() < 2e-5
    seeded = np.load(os.path.join(os.path.dirname(TRAINED), "bundled_calls_seed7.npz"))
    lab = np.array([[int(c) for c in str(s)[:150].ljust(150, "0")] for s in g.labels])[:, 2:148]
    assert (z["calls"][:, 2:148] == lab).mean() > (seeded["calls"][:, 2:148] == lab).mean()
