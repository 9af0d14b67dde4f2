"""The C-ABI library builds, loads and exports every symbol include/deepore_b200.h declares (no GPU needed)."""
import ctypes


def test_library_exports_all_declared_symbols():
    from deepore_b200 import _lib
    L = _lib.lib()
    names = _lib.exported_symbols()
    assert "dp_encode_windows" in names and "dp_last_error" in names
    for n in names:
        assert hasattr(L, n), "missing export %s" % n
    assert L.dp_version() >= 100


def test_no_cpu_fallback_without_gpu():
    import pytest
    import torch
    from deepore_b200 import _lib
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(_lib.DeeporeError):
        _lib.require_cuda()
