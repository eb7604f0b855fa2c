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
    assert lib.roma_match_keypoints_static_max_points() == 16384
    # tile map of the shared-patch local-correlation kernel (one byte per 8 x 8 tile); full correlation for r = 7; nothing otherwise
    assert lib.roma_refiner_gather_workspace_bytes(16, 108, 108, 3) >= 16 * 14 * 14
    assert lib.roma_refiner_gather_workspace_bytes(16, 40, 40, 7) >= 16 * 1600 * 1600 * 4
    assert lib.roma_refiner_gather_workspace_bytes(16, 432, 432, 0) == 0 and lib.roma_refiner_gather_workspace_bytes(2, 100, 100, 7) == 0


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


def test_late_entry_points_validate_and_refuse_cpu():
    """The input-path, GELU and LayerNorm-shift entry points: argument checks before any CUDA call; wrappers refuse
    CPU tensors (there is no CPU path); the kde workspace covers the column partials of the symmetric evaluation."""
    import torch
    from roma_b200 import _lib, ops
    lib = _lib.load()
    assert lib.roma_prepare_images_pil(None, None, None, 1, 8, 8, 4, 4, 192, 64, 8, 1, None, None, None) == -1
    assert b"null" in lib.roma_last_error_string()
    assert lib.roma_prepare_images(None, None, 1, 8, 8, 4, 4, None, None, None) == -1
    assert lib.roma_gelu(None, None, 8, 1, None) == -1
    assert lib.roma_layernorm_shift(None, None, None, None, None, 1, 8, 1e-6, 1, None) == -1
    N = 40000
    assert lib.roma_kde_workspace_bytes(N) >= 64 * N * 4 + ((N + 127) // 128) * N * 4
    assert lib.roma_kde_workspace_bytes(N) >= N * 512 * 4 + N * 32 + 32768 * 4          # exact mode: per-thread row sums + packed operands + table
    assert lib.roma_kde_workspace_bytes(200000) < 200000 * 512 * 4 + (16 << 20)        # beyond the symmetric limit: no column partials
    for call in (lambda: ops.gelu(torch.zeros(16, dtype=torch.bfloat16)),
                 lambda: ops.prepare_images(torch.zeros(1, 3, 8, 8, dtype=torch.uint8), (4, 4)),
                 lambda: ops.prepare_images_pil(torch.zeros(1, 3, 8, 8, dtype=torch.uint8), (4, 4)),
                 lambda: ops.layer_norm(torch.zeros(2, 8), torch.ones(8), torch.zeros(8), 1e-6, shift=torch.zeros(8))):
        with pytest.raises(RuntimeError, match="no CPU fallback"):
            call()
