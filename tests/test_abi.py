"""The C-ABI library loads and exports every symbol include/roma_b200.h declares (no compute, no GPU)."""
import re
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent


def declared_symbols():
    text = (ROOT / "include" / "roma_b200.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(roma_[a-z0-9_]+)\s*\(", text)))


def test_header_and_binding_agree():
    from roma_b200 import _lib
    assert declared_symbols() == sorted(_lib.SIGNATURES), "include/roma_b200.h and roma_b200/_lib.py list different entry points"


def test_library_loads_and_exports_everything():
    from roma_b200 import _lib, build
    build.build()
    lib = _lib.load()
    assert lib.roma_abi_version() == _lib.ABI_VERSION
    for name in declared_symbols():
        assert hasattr(lib, name), name
    assert lib.roma_gp_workspace_bytes(2, 1600, 1600, 512, 512) > 2 * 1600 * 1600 * 4
    assert lib.roma_kde_workspace_bytes(1000) >= 4000
    assert lib.roma_match_keypoints_workspace_bytes(10, 10) > 0


def test_bad_arguments_are_rejected_without_a_gpu():
    """Argument validation happens before any CUDA call, so it is testable on the CPU box."""
    import ctypes as C
    from roma_b200 import _lib
    lib = _lib.load()
    s = _lib.Strides4(0, 0, 0, 0)
    rc = lib.roma_refiner_gather(None, None, None, s, None, None, 0.0, None, 1, 0, -1, -1, -1, 1, 1, 1, 1, 1, 1, 0, 0, 0, None)
    assert rc == -1 and b"null" in lib.roma_last_error_string()
    rc = lib.roma_gp_posterior(None, None, None, None, None, 1, 1, 1, 1, 1, 1, 1, 0.2, 0.1, 1e-6, 0, None, 0, None)
    assert rc == -1
    with pytest.raises(ValueError):
        _lib.check(rc, "roma_gp_posterior")


def test_ops_refuse_cpu_tensors():
    import torch
    from roma_b200 import ops
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ops.kde(torch.zeros(10, 4))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ops.gp_posterior(torch.zeros(1, 8, 2, 2), torch.zeros(1, 8, 2, 2), torch.zeros(4, 2), torch.zeros(4))
