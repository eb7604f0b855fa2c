"""GPU: roma_gp_posterior with the finished panels kept as bf16 hi / lo planes and read through TMA (roma_gp_set_planes(1))
against the on-the-fly split form (same products, another accumulation order) and the fp64 oracle."""
import pytest
import torch

from oracle import roma_oracle as O

pytestmark = pytest.mark.gpu
PLANES_DEFAULT = 1


@pytest.mark.parametrize("path", [2, 3], ids=["3-plane", "2-plane"])
@pytest.mark.parametrize("side_x,side_y,B,dtype", [(40, 40, 4, torch.bfloat16), (24, 20, 2, torch.float32), (48, 48, 2, torch.bfloat16),
                                                  (12, 12, 3, torch.float16)])
def test_gp_planes_equals_split_on_the_fly(cuda, side_x, side_y, B, dtype, path):
    from roma_b200 import _lib, ops
    lib = _lib.load()
    g = torch.Generator().manual_seed(side_x + side_y)
    C = D = 512
    base = O.bilinear_resize(torch.randn(B, C, 6, 6, generator=g), (side_y, side_y))
    y = (base + 0.3 * torch.randn(B, C, side_y, side_y, generator=g)).to(dtype)
    x = (O.bilinear_resize(base, (side_x, side_x)) + 0.3 * torch.randn(B, C, side_x, side_x, generator=g)).to(dtype)
    pw, pb = torch.randn(D, 2, generator=g) * 0.5, torch.randn(D, generator=g) * 0.5
    cl = lambda t: t.to(cuda).contiguous(memory_format=torch.channels_last)
    try:
        lib.roma_gp_set_path(path)                                           # 2: 3-plane solve (the default), 3: 2-plane
        lib.roma_gp_set_planes(0)
        mu0 = ops.gp_posterior(cl(x), cl(y), pw.to(cuda), pb.to(cuda))
        lib.roma_gp_set_planes(1)
        n0 = lib.roma_kernel_launches()
        mu1 = ops.gp_posterior(cl(x), cl(y), pw.to(cuda), pb.to(cuda))
        mu2 = ops.gp_posterior(cl(x), cl(y), pw.to(cuda), pb.to(cuda))
    finally:
        lib.roma_gp_set_planes(PLANES_DEFAULT)
        lib.roma_gp_set_path(0)
    want = O.gp_posterior(x[:1].float(), y[:1].float(), pw, pb, dtype=torch.float64)
    e0 = (mu0[:1].cpu().double() - want).abs().max().item()
    e1 = (mu1[:1].cpu().double() - want).abs().max().item()
    print(f"GP {side_x}^2 x {side_y}^2 {dtype}: on-the-fly split {e0:.2e}, planes {e1:.2e} vs fp64; planes vs split {(mu1 - mu0).abs().max().item():.2e}")
    assert torch.equal(mu1, mu2)                                         # deterministic
    assert e1 < 1e-3 and e1 < 3 * e0 + 1e-5
    assert (mu1 - mu0).abs().max().item() < (2e-4 if path == 3 else 2e-5)
