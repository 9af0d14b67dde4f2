"""CPU tests of the drop-in boundary: module names, constructor/state_dict parity with the reference Model, and
unpickling of whole-object checkpoints written by the reference's own class (needs /root/reference, build container)."""
import io
import os
import sys

import numpy as np
import pytest

torch = pytest.importorskip("torch")
REF = "/root/reference/src/modules/model.py"


def test_model_has_reference_surface():
    from deepore_b200.model import Model
    m = Model()                                        # reference defaults (model.py:60)
    assert (m.input_channel, m.output_channel, m.coverage_depth, m.hidden_size, m.hidden_layer, m.classN) == (3, 256, 50, 500, 3, 3)
    m = Model(input_channel=4, coverage_depth=200)
    keys = list(m.state_dict().keys())
    assert len(keys) == 17 and not any(k.startswith("rnns") for k in keys)      # GRUs are unregistered
    assert sum(v.numel() for k, v in m.state_dict().items() if not k.endswith("num_batches_tracked")) == 26262995      # same as the reference class (checked against it in the build container)
    assert [n for n, _ in m.rnns] == ["0", "1", "2"]
    assert m.init_hidden(5).shape == (6, 5, 500)
    with pytest.raises(Exception):
        m(torch.zeros(1, 4, 3, 200), m.init_hidden(1))  # CPU tensors: no fallback


def test_compat_names_resolve():
    import deepore_b200.compat as c
    c.install()
    import importlib
    mm = importlib.import_module("modules.model")
    md = importlib.import_module("modules.dataset")
    from deepore_b200.model import Model
    assert mm.Model is Model and hasattr(md, "PileupDataset") and hasattr(md, "TextColor")
    import torchnet.meter as meter
    meter.ConfusionMeter(3)


@pytest.mark.skipif(not os.path.isfile(REF), reason="needs /root/reference")
def test_reference_pickle_loads_into_our_class(tmp_path):
    """torch.save(model) by the REFERENCE class (train.py:194) -> torch.load under the compat names gives our class
    with identical tensors (registered ones and the GRUs of the unregistered `rnns` list)."""
    import subprocess
    from deepore_b200 import modelio
    pkl = str(tmp_path / "ref_model.pkl")
    code = (
        "import sys, torch; sys.path.insert(0, '/root/reference/src'); sys.path.insert(0, %r)\n"
        "from modules.model import Model\n"
        "from deepore_b200 import modelio\n"
        "m = Model(input_channel=4, coverage_depth=200); modelio.load_into_module(m, modelio.seeded_state(3)); m.eval()\n"
        "torch.save(m, %r)\n" % (os.path.dirname(os.path.dirname(os.path.abspath(__file__))), pkl))
    subprocess.check_call([sys.executable, "-c", code])
    import deepore_b200.compat as c
    c.install()
    for k in [k for k in sys.modules if k == "modules" or k.startswith("modules.")]:
        del sys.modules[k]
    m = torch.load(pkl)
    from deepore_b200.model import Model
    assert type(m) is Model
    st, want = modelio.state_from_module(m), modelio.seeded_state(3)
    for k, v in want.items():
        assert np.array_equal(np.asarray(st[k]), v), k


def test_torch_load_patch_is_scoped_and_reversible():
    """compat.install() warns that torch.load trusts whole-object pickles from then on; uninstall() and the
    trusted_checkpoints() context manager put the stock torch.load back."""
    import warnings
    import deepore_b200.compat as c
    c.uninstall()
    stock = torch.load
    with c.trusted_checkpoints():
        assert getattr(torch.load, "_deepore_patched", False)
    assert torch.load is stock
    with warnings.catch_warnings(record=True) as w:
        warnings.simplefilter("always")
        c.install()
    assert getattr(torch.load, "_deepore_patched", False) and any("weights_only" in str(x.message) for x in w)
    c.uninstall()
    assert torch.load is stock
    c.install(patch_torch_load=False)
    assert torch.load is stock
