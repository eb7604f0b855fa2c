"""patch(model) rebinding, checked on the CPU with stand-in modules (the reference itself is exercised in the build
container when /root/reference is present)."""
import sys
import tempfile
import types
from pathlib import Path

import pytest
import torch
import torch.nn as nn

import roma_b200
from roma_b200 import ops


def _standins():
    m, t = types.ModuleType("matcher_standin"), types.ModuleType("tiny_standin")
    m.local_correlation = m.kde = m.cls_to_flow_refine = t.kde = None

    class GP(nn.Module):
        def __init__(self):
            super().__init__()
            self.pos_conv = nn.Conv2d(2, 8, 1)
            self.K = types.SimpleNamespace(T=0.2)
            self.sigma_noise = 0.1

        def forward(self, x, y):
            return "reference"

    model = nn.Module()
    model.decoder = nn.Module()
    model.decoder.gps = nn.ModuleDict({"16": GP()})
    return model, m, t


def test_patch_rebinds_all_seams_and_has_no_cpu_fallback():
    model, m, t = _standins()
    assert roma_b200.patch(model, m, t) is model
    assert m.local_correlation is ops.local_correlation and m.kde is ops.kde and t.kde is ops.kde
    x = torch.zeros(1, 8, 4, 4)
    with pytest.raises(RuntimeError):                       # CUDA only: fails loudly on CPU tensors
        model.decoder.gps["16"](x, x)
    with pytest.raises(RuntimeError):
        m.local_correlation(x, x, 2, flow=torch.zeros(1, 2, 4, 4))
    with pytest.raises(RuntimeError):
        m.cls_to_flow_refine(torch.zeros(1, 16, 4, 4))


@pytest.mark.skipif(not Path("/root/reference/romatch").exists(), reason="reference only exists in the build container")
def test_patch_real_reference_model():
    stub = Path(tempfile.mkdtemp(prefix="kornia_stub_")); (stub / "kornia").mkdir(); (stub / "kornia" / "__init__.py").write_text("")
    sys.path[:0] = [str(stub), "/root/reference"]
    try:
        import romatch.models.matcher as M
        import romatch.models.tiny as T
        gp = M.GP(M.CosKernel, T=0.2, gp_dim=16, basis="fourier", no_cov=True)
        model = nn.Module(); model.decoder = nn.Module(); model.decoder.gps = nn.ModuleDict({"16": gp})
        saved = (M.local_correlation, M.kde, M.cls_to_flow_refine, T.kde)
        try:
            roma_b200.patch(model)
            assert M.local_correlation is ops.local_correlation and T.kde is ops.kde
            with pytest.raises(RuntimeError):
                gp(torch.zeros(1, 16, 4, 4), torch.zeros(1, 16, 4, 4))
        finally:
            M.local_correlation, M.kde, M.cls_to_flow_refine, T.kde = saved
    finally:
        del sys.path[:2]
