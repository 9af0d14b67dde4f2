"""modules.dataset drop-in (deepore_b200/dataset.py) against the reference's own class
(/root/reference/src/modules/dataset.py:26-57): same items for the same CSV -- float32 (4, 150, 200) = u8 / 255 after
ToTensor, int64 label vector -- and the in-memory variant that serves encoder tensors without the PNG round trip.

Two facts about the reference class that the tests pin down (both checked against it in the build container):
  * it reads the CSV without dtype=str, so pandas parses a label as ONE big integer: digits survive only because every label
    the generator writes starts with a class digit 1..3 (a leading '0' would be dropped).  The drop-in reads the column as text;
  * it stacks the label vectors with np.array(), which needs equal lengths on numpy >= 1.24 (labels grow with the insert rows
    of a window, Pileup.py:320-333); the drop-in keeps ragged labels as an object array.
"""
import importlib.util
import os
import sys

import numpy as np
import pytest

import golden_util as gu

torch = pytest.importorskip("torch")
REF_DATASET = "/root/reference/src/modules/dataset.py"


def _write_fixture(tmp_path, pick, labels):
    from PIL import Image
    g = gu.load("bundled_na12878")
    lines = []
    for k, lab in zip(pick, labels):
        p = str(tmp_path / ("w%d.png" % k))
        Image.fromarray(np.ascontiguousarray(g.img[k]), "RGBA").save(p, "PNG")
        lines.append("%s,%s\n" % (p, lab))
    csv = str(tmp_path / "summary-3.csv")
    open(csv, "w").writelines(lines)
    return g, csv


def _to_tensor():
    from torchvision import transforms
    return transforms.Compose([transforms.ToTensor()])           # test.py:28


def test_items_are_totensor_of_the_png_and_int64_labels(tmp_path):
    from deepore_b200.dataset import InMemoryPileupDataset, PileupDataset
    gold = np.load(os.path.join(gu.GOLD, "bundled_windows.npz"))
    pick = [0, 12, 57, 20, 259]
    labels = [str(gold["label_out"][k]) for k in pick]          # ragged: 201 .. 212 digits
    assert len(set(len(x) for x in labels)) > 1
    g, csv = _write_fixture(tmp_path, pick, labels)
    ds = PileupDataset(csv, _to_tensor())
    assert len(ds) == len(pick)
    for i, k in enumerate(pick):
        img, lab = ds[i]
        assert img.dtype == torch.float32 and tuple(img.shape) == (4, 150, 200)
        want = torch.from_numpy(g.img[k]).permute(2, 0, 1).float() / 255.0
        assert torch.equal(img, want)
        assert lab.dtype == torch.int64 and lab.tolist() == [int(c) for c in labels[i]]
    # the in-memory dataset serves the same items straight from the encoder tensors, through a DataLoader as well
    mem = InMemoryPileupDataset(g.img[pick], labels)
    assert len(mem) == len(pick)
    for i in range(len(pick)):
        a, b = ds[i], mem[i]
        assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1])
    from torch.utils.data import DataLoader
    clip = lambda batch: (torch.stack([x for x, _ in batch]), torch.stack([y[:150] for _, y in batch]))   # noqa: E731
    for (ia, la), (ib, lb) in zip(DataLoader(ds, batch_size=2, shuffle=False, collate_fn=clip),
                                  DataLoader(mem, batch_size=2, shuffle=False, collate_fn=clip)):
        assert torch.equal(ia, ib) and torch.equal(la, lb) and ia.shape[1:] == (4, 150, 200)


def test_missing_png_is_refused(tmp_path):
    from deepore_b200.dataset import PileupDataset
    csv = str(tmp_path / "s.csv")
    open(csv, "w").write("%s,111\n" % (tmp_path / "nope.png"))
    with pytest.raises(AssertionError):                          # dataset.py:34-35
        PileupDataset(csv)


@pytest.mark.skipif(not os.path.isfile(REF_DATASET), reason="needs /root/reference (build container)")
def test_same_items_as_the_reference_class(tmp_path):
    """The UNMODIFIED reference class on the same CSV (labels of equal length, as its np.array() stacking needs)."""
    if not hasattr(np, "int"):
        np.int = int                                             # dataset.py:44 predates numpy 1.24
    spec = importlib.util.spec_from_file_location("ref_dataset", REF_DATASET)
    ref = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ref)
    from deepore_b200.dataset import PileupDataset
    gold = np.load(os.path.join(gu.GOLD, "bundled_windows.npz"))
    pick = [k for k in range(260) if len(str(gold["label_out"][k])) == 201][:6]
    labels = [str(gold["label_out"][k]) for k in pick]
    _, csv = _write_fixture(tmp_path, pick, labels)
    a, b = ref.PileupDataset(csv, _to_tensor()), PileupDataset(csv, _to_tensor())
    assert len(a) == len(b) == len(pick)
    for i in range(len(pick)):
        (ia, la), (ib, lb) = a[i], b[i]
        assert ia.dtype == ib.dtype and torch.equal(ia, ib)
        assert la.dtype == lb.dtype == torch.int64 and torch.equal(la, lb)
    assert ref.TextColor.PURPLE == sys.modules["deepore_b200.dataset"].TextColor.PURPLE
