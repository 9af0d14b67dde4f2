"""The reference's inference driver (/root/reference/src/test.py:27-111) run against the drop-in, statement by statement.

`/root/reference` does not exist on the GPU box, so the loop of test.py:51-111 is restated here (as
oracle/make_calls_golden.py restates it for the unmodified reference model on the CPU) over the very objects test.py
would get through the compat module names:

    windows.generatePileupBasedonVCF  ->  PNG per window + summary-3.csv            (pileupGenerator.py:97-123)
    modules.dataset.PileupDataset(csv, ToTensor) + DataLoader(batch_size=100)         (test.py:28-38)
    torch.load(whole-object pickle) -> modules.model.Model, .cuda(), .eval()          (test.py:42-45)
    init_hidden, images[:, :, row:row+3, :], model(x, hidden), topk(1), most_common, confusion_tensor[t][p] += 1

Two things the committed reference cannot do by itself on these labels, and how they are handled (both on the label side,
neither touches the model path): (1) label vectors grow with the insert rows of a window (201 .. 212 digits) and the
default collate cannot stack them -- the loop only ever reads `pl[:, row]` for row < 150, so the collate keeps the first 150
digits; (2) Pileup.py labels rows with FOUR symbols (0 none/insert, 1 hom-ref, 2 het, 3 hom-alt) while test.py's
confusion tensor is 3 x 3 (test.py:49) -- class 3 is clipped to 2, as any 3-class run of that script has to.

The expected confusion tensor comes from the UNMODIFIED reference model's votes (tests/golden/bundled_calls_seed7.npz) and
the unmodified generator's labels (tests/golden/bundled_windows.npz).
"""
import os
import sys

import numpy as np
import pytest

import golden_util as gu

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

VCF = os.path.join(gu.GOLD, "bundled_confident.vcf")


def most_common(lst):                                   # test.py:23-24 (ints instead of 0-d tensors: a set of tensors
    return max(sorted(set(lst)), key=lst.count)         # hashes by identity; sorted() fixes the 3-way tie to the smallest)


def _golden_confusion():
    gc = np.load(os.path.join(gu.GOLD, "bundled_calls_seed7.npz"))
    gw = np.load(os.path.join(gu.GOLD, "bundled_windows.npz"))
    conf = np.zeros((3, 3), np.int64)
    for i in range(260):
        lab = [min(int(c), 2) for c in str(gw["label_out"][i])[:150]]
        for row in range(2, 148):
            conf[lab[row], int(gc["calls"][i, row])] += 1
    return conf, gc


@pytest.mark.parametrize("precision", ["fp16x3", "fp16"])
def test_reference_test_loop_against_the_drop_in(tmp_path, precision):
    from torch.utils.data import DataLoader
    from torchvision import transforms
    import deepore_b200.compat as compat
    from deepore_b200 import model as dmodel, modelio, windows
    compat.install()
    import importlib
    PileupDataset = importlib.import_module("modules.dataset").PileupDataset
    # ---- pileupGenerator.py: PNG + CSV for the 260 qualifying records of the bundled VCF
    g = gu.load("bundled_na12878")
    bam, fa = str(tmp_path / "b.bam"), str(tmp_path / "r.fa")
    gu.write_bam(bam, g.reads)
    gu.write_fasta(fa, g.ref, g.ref_base)
    out = str(tmp_path) + "/"
    windows.allVariantRecord.clear()
    try:
        assert windows.generatePileupBasedonVCF("3", bam, fa, VCF, out, 100) == 260
    finally:
        windows.allVariantRecord.clear()
    csvFile = out + "summary-3.csv"
    # ---- a whole-object checkpoint, as train.py:194 saves it and test.py:42 loads it
    pkl = str(tmp_path / "CNN.pkl")
    m0 = dmodel.from_state(modelio.seeded_state(7), "cpu")
    m0.set_precision(precision)
    torch.save(m0, pkl)
    del m0

    # ---- test.py:27-111, restated
    batchSize, gpu_mode = 100, True
    transformations = transforms.Compose([transforms.ToTensor()])
    test_dset = PileupDataset(csvFile, transformations)

    def collate(batch):                                  # see the module docstring, (1)
        return torch.stack([x for x, _ in batch]), torch.stack([y[:150] for _, y in batch])
    testloader = DataLoader(test_dset, batch_size=batchSize, shuffle=False, num_workers=0, pin_memory=gpu_mode,
                            collate_fn=collate)
    model = torch.load(pkl)                              # compat.install() made weights_only=False the default
    assert type(model).__module__ in ("deepore_b200.model", "modules.model")
    if gpu_mode:
        model = model.cuda()
    model.eval()
    seq_len = 3
    confusion_tensor = torch.zeros(3, 3)
    n_images = 0
    for counter, (images, labels) in enumerate(testloader):
        hidden = model.init_hidden(images.size(0))
        pl = labels
        if gpu_mode:
            images = images.cuda()
        prediction_stack = []
        for row in range(0, images.size(2), 1):
            if row + seq_len > images.size(2):
                continue
            x = images[:, :, row:row + seq_len, :]
            ypl = pl[:, row]
            preds = model(x, hidden)
            assert tuple(preds.shape) == (images.size(0), seq_len, 3)
            preds = preds.data.topk(1)[1].cpu()
            prediction_stack.append(preds)
            if row + 1 >= seq_len:
                for i in range(images.size(0)):
                    pr = []
                    k = seq_len - 1
                    for j in range(len(prediction_stack)):
                        pr.append(int(prediction_stack[j][i][k][0]))
                        k -= 1
                    p = most_common(pr)
                    t = min(int(ypl[i]), 2)              # see the module docstring, (2)
                    confusion_tensor[t][p] += 1
                prediction_stack.pop(0)
        n_images += images.size(0)
    assert n_images == 260
    want, gc = _golden_confusion()
    got = confusion_tensor.numpy().astype(np.int64)
    assert got.sum() == want.sum() == 260 * 146
    moved = int(np.abs(got - want).sum())
    print("\n[test.py loop, %s] confusion tensor\n%s\nreference model's\n%s\nentries moved: %d" % (precision, got, want, moved))
    if precision == "fp16x3":
        assert moved == 0, "the split-precision mode must reproduce the reference's confusion tensor exactly"
    else:
        assert moved <= 2 * int(0.003 * 260 * 146)      # each differing vote (counted in test_model_gpu.py) moves two entries


@pytest.mark.parametrize("precision,tol", [("fp16x3", 2e-5), ("fp16", 2e-3)])
def test_validation_pass_of_train_py(precision, tol):
    """The validation pass of the reference's train.py (:21-78: forward of every 3-row call + nn.CrossEntropyLoss on
    outputs.view(-1, 3) vs labels[:, row:row+3], loss summed over the calls, divided by calls x batch) on the CUDA forward
    (Model.validation_loss -> dp_cross_entropy), against the same quantity computed with torch's own cross_entropy from the
    UNMODIFIED reference model's logits (tests/golden/bundled_calls_seed7.npz).  Deterministic part only: the reference skips
    most all-reference calls by drawing random numbers (train.py:62-67).  Labels: first 150 digits, class 3 clipped to 2."""
    import torch.nn.functional as F
    from deepore_b200 import model as dmodel, modelio
    gc = np.load(os.path.join(gu.GOLD, "bundled_calls_seed7.npz"))
    gw = np.load(os.path.join(gu.GOLD, "bundled_windows.npz"))
    g = gu.load("bundled_na12878")
    N = 64
    labels = np.array([[min(int(c), 2) for c in str(gw["label_out"][i])[:150]] for i in range(N)], np.int64)
    ref_logits = torch.from_numpy(gc["logits"][:N])                                # [N, 148, 3, 3]
    want_calls = []
    for row in range(148):
        y = torch.from_numpy(labels[:, row:row + 3])
        want_calls.append(float(F.cross_entropy(ref_logits[:, row].reshape(-1, 3), y.reshape(-1))))
    want_total = sum(want_calls) / (148 * N)
    m = dmodel.from_state(modelio.seeded_state(int(gc["seed"])), "cuda", precision)
    per_call, total = m.validation_loss(torch.from_numpy(g.img[:N]).cuda(), torch.from_numpy(labels).cuda())
    got = per_call.cpu().numpy()
    assert got.shape == (148,)
    assert np.abs(got - np.asarray(want_calls)).max() < tol and abs(total - want_total) < tol
    # the kernel by itself against torch on the same logits, and its refusal of a label without a class
    lg = ref_logits.cuda()
    y = torch.from_numpy(np.stack([labels[:, r:r + 3] for r in range(148)], axis=1)).cuda()      # [N, 148, 3]
    loss, s, bad = dmodel.cross_entropy(lg, y)
    ref = F.cross_entropy(lg.reshape(-1, 3), y.reshape(-1), reduction="none").reshape(N, 148, 3)
    assert bad == 0 and torch.allclose(loss, ref, atol=2e-6) and abs(s - float(ref.double().sum())) < 1e-3
    y[0, 0, 0] = 3
    loss, s, bad = dmodel.cross_entropy(lg, y)
    assert bad == 1 and torch.isinf(loss[0, 0, 0])
    with pytest.raises(Exception):
        m.validation_loss(torch.from_numpy(g.img[:2]).cuda(), torch.full((2, 150), 3))
