"""CPU-only checks of the drop-in boundary: libqb200.so loads, exports every symbol that
include/qb200.h declares, _abi.py binds every one of them, and - without a GPU - the library fails
loudly instead of falling back to anything."""
import ctypes
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_functions():
    text = open(os.path.join(ROOT, "include", "qb200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(qb_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_exported_and_bound():
    import qtorch_b200  # noqa: F401
    from qtorch_b200 import _abi
    names = declared_functions()
    assert len(names) >= 18
    lib = ctypes.CDLL(_abi.LIB_PATH)
    for n in names:
        assert hasattr(lib, n), "libqb200.so does not export %s" % n
        assert n in _abi.PROTOTYPES, "_abi.py has no prototype for %s" % n
    assert sorted(_abi.PROTOTYPES) == names
    assert _abi.load().qb_version() >= 100


def test_ws_size_helpers_need_no_gpu():
    import qtorch_b200  # noqa: F401
    from qtorch_b200 import _abi
    lib = _abi.load()
    assert lib.qb_minmax_ws_bytes() > 0
    assert lib.qb_measure_ws_bytes(256) >= 256 * 8 + 64
    assert lib.qb_rangebn_ws_bytes(64, 16) == 64 * 16 * 16
    assert lib.qb_rangebn_bwd_ws_bytes(64) > 0


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_no_cpu_fallback():
    import qtorch_b200  # noqa: F401
    from qtorch_b200 import _abi
    with pytest.raises(_abi.QB200Error):
        _abi.device_info(0)
    with pytest.raises(_abi.QB200Error):
        _abi.fakequant_affine(torch.zeros(8), 8, lo=0.0, hi=1.0)


def test_argument_errors_reported():
    import qtorch_b200  # noqa: F401
    from qtorch_b200 import _abi
    lib = _abi.load()
    # argument validation happens before any CUDA call
    rc = lib.qb_fakequant_affine(None, None, None, 16, 8, None, 0.0, 1.0, 0, None)
    assert rc == -1 and "null" in _abi.last_error()
    rc = lib.qb_fakequant_affine(1, 1, None, 16, 99, None, 0.0, 1.0, 0, None)
    assert rc == -1 and "bits" in _abi.last_error()
    rc = lib.qb_rangebn_stats(1, 3, 2, 5, 16, None, None, 1, 1, 1 << 20, None)
    assert rc == -1 and "divisible" in _abi.last_error()


def test_every_entry_point_is_documented():
    """INTEGRATION.md names every function include/qb200.h declares (the reference-side binding table stays complete)."""
    text = open(os.path.join(ROOT, "INTEGRATION.md")).read()
    missing = [n for n in declared_functions() if n not in text
               and not (n.endswith("_supported") and "`*_supported` functions" in text)]
    assert not missing, missing
