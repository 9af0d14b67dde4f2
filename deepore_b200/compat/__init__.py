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
"""
import os
import sys
import types

HERE = os.path.dirname(os.path.abspath(__file__))


def install():
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
    import torch
    if not getattr(torch.load, "_deepore_patched", False):
        _orig = torch.load

        def load(*a, **k):
            k.setdefault("weights_only", False)         # whole-object pickles (train.py:194 / test.py:42)
            return _orig(*a, **k)

        load._deepore_patched = True
        torch.load = load
    return HERE
