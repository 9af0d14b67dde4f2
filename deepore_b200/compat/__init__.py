"""Import-name compatibility for running the reference's own drivers against deepore_b200.

The reference's scripts import `modules.model`, `modules.dataset` (src/test.py:10, src/train.py:15-16) and unpickle
whole `modules.model.Model` objects (src/test.py:42).  Putting THIS directory on sys.path (``install()`` or
``PYTHONPATH=deepore_b200/compat``) makes those names resolve to the B200 implementations, so `src/test.py` runs
unchanged:

    python -c "import deepore_b200.compat as c; c.install(); import runpy, sys; \
               sys.argv=['test.py','--test_file',CSV,'--model_path',PKL,'--gpu_mode','1']; \
               runpy.run_path('/root/reference/src/test.py', run_name='__main__')"

install() also provides what the 2017-era script expects from its environment and this image lacks or changed:
`torchnet.meter.ConfusionMeter` (constructed at test.py:47, never used), `scipy.misc` (imported, never used),
`np.int` (dataset.py:44) and `torch.load(..., weights_only=False)` as the default (test.py:42 loads a whole object).
Nothing here touches the reference's files.

TRUST: a whole-object checkpoint is a pickle, and unpickling executes code chosen by whoever wrote the file.  The patched
`torch.load` default is process-wide, so after install() EVERY torch.load of the process trusts its file the way the
reference's test.py does; install() says so in a warning.  Load only checkpoints you wrote; `uninstall()` restores
torch.load, and `trusted_checkpoints()` scopes the patch to a `with` block (pass patch_torch_load=False to install()).
"""
import os
import sys
import types

HERE = os.path.dirname(os.path.abspath(__file__))


_ORIG_LOAD = None


def _patch_torch_load():
    global _ORIG_LOAD
    import torch
    if getattr(torch.load, "_deepore_patched", False):
        return
    _ORIG_LOAD = torch.load
    orig = torch.load

    def load(*a, **k):
        k.setdefault("weights_only", False)             # whole-object pickles (train.py:194 / test.py:42)
        return orig(*a, **k)

    load._deepore_patched = True
    torch.load = load


def uninstall():
    """Restore torch.load (the module names stay importable)."""
    global _ORIG_LOAD
    import torch
    if _ORIG_LOAD is not None and getattr(torch.load, "_deepore_patched", False):
        torch.load = _ORIG_LOAD
    _ORIG_LOAD = None


class trusted_checkpoints:
    """`with trusted_checkpoints(): model = torch.load(path)` -- torch.load unpickles whole objects only inside the block."""

    def __enter__(self):
        self._was = _ORIG_LOAD is not None
        _patch_torch_load()
        return self

    def __exit__(self, *exc):
        if not self._was:
            uninstall()
        return False


def install(patch_torch_load=True):
    if HERE not in sys.path:
        sys.path.insert(0, HERE)
    import numpy as np
    if not hasattr(np, "int"):
        np.int = int                                    # dataset.py:44 (removed in numpy 1.24)
    try:
        import torchnet  # noqa: F401
    except Exception:
        tn = types.ModuleType("torchnet")
        meter = types.ModuleType("torchnet.meter")

        class ConfusionMeter:                           # test.py:47 constructs one and never calls it
            def __init__(self, k, normalized=False):
                self.k = k

            def add(self, *a, **k):
                pass

        meter.ConfusionMeter = ConfusionMeter
        tn.meter = meter
        sys.modules["torchnet"] = tn
        sys.modules["torchnet.meter"] = meter
    try:
        from scipy import misc  # noqa: F401
    except Exception:
        import scipy
        m = types.ModuleType("scipy.misc")
        scipy.misc = m
        sys.modules["scipy.misc"] = m
    if patch_torch_load:
        import warnings
        import torch
        if not getattr(torch.load, "_deepore_patched", False):
            warnings.warn("deepore_b200.compat.install(): torch.load now defaults to weights_only=False for this process "
                          "(the reference's test.py loads whole pickled objects); load only checkpoints you trust, "
                          "or use compat.trusted_checkpoints() / compat.uninstall()", stacklevel=2)
        _patch_torch_load()
    return HERE
