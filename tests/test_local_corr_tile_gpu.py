"""GPU: the shared-patch local-correlation kernel (local_corr_tile.cuh: 8 x 8 pixel tiles, TMA + tcgen05, coherent flows)
against the per-pixel tensor-core kernel it complements (roma_local_corr_set_path(2)) and against the fp32 reference
formulation (oracle, ATen mode).  Segments x / x_hat / emb must be bit-identical between the two paths (same arithmetic);
the correlations differ only by the fp32 summation order of the tensor core -> within 1.5 ulps of the 16-bit output."""
import math

import pytest
import torch
import torch.nn.functional as F

from oracle import roma_oracle as O

pytestmark = pytest.mark.gpu
DEFAULT_PATH = 0          # library default of roma_local_corr_set_path


@pytest.fixture(scope="module")
def ops(cuda):
    from roma_b200 import ops as _ops
    return _ops


def _grid(B, h, w, dev):
    ys = torch.linspace(-1 + 1 / h, 1 - 1 / h, h, device=dev)
    xs = torch.linspace(-1 + 1 / w, 1 - 1 / w, w, device=dev)
    return torch.stack((xs[None, :].expand(h, w), ys[:, None].expand(h, w)))[None].expand(B, 2, h, w).contiguous()


def _flow(kind, B, h, w, dev, g):
    grid = _grid(B, h, w, dev)
    if kind == "incoherent":
        return torch.rand(B, 2, h, w, device=dev, generator=g) * 2.2 - 1.1
    low = torch.randn(B, 2, max(h // 8, 2), max(w // 8, 2), device=dev, generator=g) * 0.05
    flow = grid + F.interpolate(low, size=(h, w), mode="bilinear", align_corners=False)
    if kind == "coherent":
        return flow
    if kind == "shifted":                 # coherent but pointing partly outside the image: TMA zero fill = zeros padding
        return flow + torch.tensor([0.9, -0.8], device=dev).view(1, 2, 1, 1)
    if kind == "mixed":                   # left half coherent, right half incoherent, one NaN and one Inf vector
        rnd = torch.rand(B, 2, h, w, device=dev, generator=g) * 2.2 - 1.1
        flow = torch.where((torch.arange(w, device=dev) >= w // 2).view(1, 1, 1, w), rnd, flow)
        flow[0, 0, 1, 2] = float("nan")
        flow[0, 1, 3, 5] = float("inf")
        return flow
    if kind == "stretch":                 # displacement gradient 0.5 px / px: tiles at the edge of the patch slack
        return grid * 1.5
    raise ValueError(kind)


def _both_paths(ops, x, y, flow, dw, db, r):
    from roma_b200 import _lib
    lib = _lib.load()
    try:
        lib.roma_local_corr_set_path(0)
        n0 = lib.roma_kernel_launches()
        d_tile = ops.refiner_inputs(x, y, flow, dw, db, 1.25, r, pad_to=8)
        # r = 2, 3 at C = 256 / 512: verdict kernel (tile map + work list) + persistent tile kernel + per-pixel kernel;
        # r = 7: full-correlation GEMM + per-pixel gather
        assert lib.roma_kernel_launches() - n0 == (2 if r == 7 else 3 if x.shape[1] in (256, 512) else 1)
        lib.roma_local_corr_set_path(2)
        d_px = ops.refiner_inputs(x, y, flow, dw, db, 1.25, r, pad_to=8)
    finally:
        lib.roma_local_corr_set_path(DEFAULT_PATH)
    return d_tile, d_px


@pytest.mark.parametrize("kind", ["coherent", "shifted", "mixed", "stretch", "incoherent"])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16], ids=["bf16", "fp16"])
@pytest.mark.parametrize("C,r,h,w,E", [(512, 3, 108, 108, 64), (256, 2, 54, 75, 32), (64, 3, 21, 37, 0), (128, 2, 8, 8, 16),
                                       (512, 7, 40, 40, 128), (64, 7, 13, 19, 8)])
def test_tile_kernel_vs_per_pixel_kernel_and_oracle(ops, cuda, C, r, h, w, E, dtype, kind):
    B = 2
    g = torch.Generator(device="cuda").manual_seed(C + r + h)
    x = torch.randn(B, h, w, C, device=cuda, generator=g).to(dtype).permute(0, 3, 1, 2)
    y = torch.randn(B, h, w, C, device=cuda, generator=g).to(dtype).permute(0, 3, 1, 2)
    flow = _flow(kind, B, h, w, cuda, g)
    dw = torch.randn(E, 2, 1, 1, device=cuda, generator=g) if E else None
    db = torch.randn(E, device=cuda, generator=g) if E else None
    d_tile, d_px = _both_paths(ops, x, y, flow, dw, db, r)
    K = (2 * r + 1) ** 2
    finite = torch.isfinite(flow).all(1, keepdim=True)
    assert torch.equal(d_tile[:, :C], d_px[:, :C]) and torch.equal(d_tile[:, :C], x)
    a, b_ = d_tile[:, C:2 * C + E], d_px[:, C:2 * C + E]
    assert torch.equal(torch.where(finite, a, torch.zeros_like(a)), torch.where(finite, b_, torch.zeros_like(b_)))     # x_hat, emb
    ct, cp = d_tile[:, 2 * C + E:2 * C + E + K].float(), d_px[:, 2 * C + E:2 * C + E + K].float()
    ct, cp = torch.where(finite, ct, torch.zeros_like(ct)), torch.where(finite, cp, torch.zeros_like(cp))
    eps = 2.0 ** -8 if dtype == torch.bfloat16 else 2.0 ** -11
    ulp = 2 * eps * torch.exp2(torch.floor(torch.log2(torch.maximum(cp.abs(), torch.full_like(cp, 2.0 ** -4)))))   # of the 16-bit output
    tol = 1.5 * ulp + 1e-6
    bad = int(((ct - cp).abs() > tol).sum().item())
    assert bad == 0, f"{bad} correlations further than 1.5 output ulps from the per-pixel kernel; max {(ct - cp).abs().max().item():.3e}"
    assert d_tile[:, 2 * C + E + K:].abs().max().item() == 0 if d_tile.shape[1] > 2 * C + E + K else True
    # and against the reference formulation in fp32 (finite flows only)
    if kind != "mixed":
        O.set_aten(True)
        try:
            want = O.local_correlation(x.float(), y.float(), r, flow=flow)
        finally:
            O.set_aten(False)
        err = (ct - want).abs()
        assert bool((err <= 0.51 * 2 * eps * want.abs() + 1e-4).all()), f"corr vs fp32 reference: {err.max().item():.3e}"


def test_tile_kernel_takes_the_coherent_tiles(ops, cuda):
    """Timing-free evidence that the shared-patch kernel did the work: with the per-pixel kernel's dot products disabled
    (path 2 on an all-coherent field equals path 0) the outputs are equal AND differ in summation order somewhere."""
    B, C, r, h, w = 2, 256, 2, 64, 64
    g = torch.Generator(device="cuda").manual_seed(1)
    x = torch.randn(B, h, w, C, device=cuda, generator=g).to(torch.bfloat16).permute(0, 3, 1, 2)
    y = torch.randn(B, h, w, C, device=cuda, generator=g).to(torch.bfloat16).permute(0, 3, 1, 2)
    flow = _flow("coherent", B, h, w, cuda, g)
    d_tile, d_px = _both_paths(ops, x, y, flow, None, None, r)
    assert not torch.equal(d_tile, d_px)                  # a different fp32 summation order shows in some last bits
    assert (d_tile.float() - d_px.float()).abs().max().item() < 0.05


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16], ids=["bf16", "fp16"])
@pytest.mark.parametrize("C,E,h,w", [(64, 16, 53, 47), (8, 0, 9, 11), (128, 8, 16, 16), (512, 64, 10, 12), (9, 6, 20, 22), (9, 6, 33, 18), (9, 6, 5, 7)])
def test_gather_only_v2_is_bit_identical_to_the_chunk_per_thread_kernel(ops, cuda, C, E, h, w, dtype):
    """r = 0 visits (scale 2, Tiny-RoMa): gather_only_v2_kernel (a thread per (pixel, x / x_hat chunk pair), packed fp32
    FMAs) against gather_only_kernel (a thread per 16-byte chunk; roma_local_corr_set_path(1) keeps it): same bits."""
    from roma_b200 import _lib
    lib = _lib.load()
    B = 3
    g = torch.Generator(device="cuda").manual_seed(C + h)
    x = torch.randn(B, h, w, C, device=cuda, generator=g).to(dtype).permute(0, 3, 1, 2)
    y = torch.randn(B, h, w, C, device=cuda, generator=g).to(dtype).permute(0, 3, 1, 2)
    dw = torch.randn(E, 2, 1, 1, device=cuda, generator=g) if E else None
    db = torch.randn(E, device=cuda, generator=g) if E else None
    for kind in ("coherent", "incoherent", "shifted"):
        flow = _flow(kind, B, h, w, cuda, g)
        try:
            lib.roma_local_corr_set_path(DEFAULT_PATH)
            d2 = ops.refiner_inputs(x, y, flow, dw, db, 1.25, None, pad_to=8)
            lib.roma_local_corr_set_path(1)
            d1 = ops.refiner_inputs(x, y, flow, dw, db, 1.25, None, pad_to=8)
        finally:
            lib.roma_local_corr_set_path(DEFAULT_PATH)
        assert torch.equal(d2, d1)
        want = F.grid_sample(y.float(), flow.permute(0, 2, 3, 1), align_corners=False, mode="bilinear")
        eps = 2.0 ** -8 if dtype == torch.bfloat16 else 2.0 ** -11
        assert bool(((d2[:, C:2 * C].float() - want).abs() <= 0.51 * 2 * eps * want.abs() + 2e-5).all())
