"""GPU: the kernels at BASELINE.json's full sizes (config 2: roma_outdoor 560 -> 864, decoder batch 16 = 8 pairs),
checked through size-independent properties — identities, linearity, impulse responses, permutation / swap
invariance, idempotence — because the CPU oracle cannot finish these sizes in seconds.  Where the oracle CAN
(one GP element at N = 1600) it is used directly.  Every call goes through the C ABI (roma_b200.ops)."""
import math

import pytest
import torch

from oracle import roma_oracle as O

pytestmark = pytest.mark.gpu

B16 = 16          # decoder batch of the 8-pair step


@pytest.fixture(scope="module")
def ops(cuda):
    from roma_b200 import _lib, ops as _ops
    assert _lib.load().roma_check_device(0) == 0
    return _ops


def _grid(B, h, w, dev):
    ys = torch.linspace(-1 + 1 / h, 1 - 1 / h, h, device=dev)
    xs = torch.linspace(-1 + 1 / w, 1 - 1 / w, w, device=dev)
    return torch.stack((xs[None, :].expand(h, w), ys[:, None].expand(h, w)))[None].expand(B, 2, h, w).contiguous()


# ------------------------------------------------------------------------------------------------ K1
def test_gp_full_size_vs_fp64_oracle(ops, cuda):
    """N = M = 1600 (40 x 40 at 560), C = D = 512, bf16 features: one element against the fp64 oracle, and all 16
    elements of a batch built from 2 distinct pairs must reproduce those 2 results (batch independence)."""
    g = torch.Generator().manual_seed(3)
    base = O.bilinear_resize(torch.randn(2, 512, 6, 6, generator=g), (40, 40))
    x = (base + 0.5 * torch.randn(2, 512, 40, 40, generator=g)).to(torch.bfloat16)
    y = (base.flip(0) * 0.7 + 0.5 * torch.randn(2, 512, 40, 40, generator=g)).to(torch.bfloat16)
    pw = torch.randn(512, 2, generator=g) * 0.5
    pb = torch.randn(512, generator=g) * 0.5
    want = O.gp_posterior(x[:1].float(), y[:1].float(), pw, pb, dtype=torch.float64)
    xb, yb = x.repeat(8, 1, 1, 1).to(cuda), y.repeat(8, 1, 1, 1).to(cuda)
    got = ops.gp_posterior(xb, yb, pw.to(cuda), pb.to(cuda))
    assert got.shape == (B16, 512, 40, 40)
    assert (got[0].float().cpu() - want[0].float()).abs().max().item() < 1e-3          # north_star budget for this stage
    for i in range(2, B16):
        assert torch.equal(got[i], got[i % 2])


# ------------------------------------------------------------------------------------------------ K2
@pytest.mark.parametrize("C,r,side", [(512, 3, 108), (256, 2, 216)])
def test_local_corr_full_size_identities(ops, cuda, C, r, side):
    """s8' / s4' visit: with the identity flow, x_hat == y exactly (bilinear weights are (1, 0) at pixel centres) and
    the centre tap of the correlation is <x, y> / sqrt(C); scaling y by 2 scales x_hat and corr by exactly 2."""
    g = torch.Generator(device="cuda").manual_seed(5)
    x = torch.randn(B16, side, side, C, device=cuda, generator=g).to(torch.bfloat16).permute(0, 3, 1, 2)
    y = torch.randn(B16, side, side, C, device=cuda, generator=g).to(torch.bfloat16).permute(0, 3, 1, 2)
    flow = _grid(B16, side, side, cuda)
    K = (2 * r + 1) ** 2
    d = ops.refiner_inputs(x, y, flow, None, None, 1.0, r, pad_to=8)
    assert d.shape[1] == (2 * C + K + 7) // 8 * 8
    assert torch.equal(d[:, :C], x)
    # identity warp: the fp32 un-normalisation of the grid leaves bilinear weights within ~1e-5 of (1, 0)
    assert (d[:, C:2 * C].float() - y.float()).abs().max().item() < 0.05               # <= one bf16 ulp at |y| ~ 5
    centre = d[:, 2 * C + K // 2].float()
    want = (x.float() * y.float()).sum(1) / math.sqrt(C)
    assert (centre - want).abs().max().item() < 2e-2 * want.abs().max().item()          # bf16 output rounding
    assert d[:, 2 * C + K:].abs().max().item() == 0                                     # pitch padding is zero
    d2 = ops.refiner_inputs(x, y * 2, flow, None, None, 1.0, r, pad_to=8)
    assert torch.equal(d2[:, C:2 * C], d[:, C:2 * C] * 2)
    assert torch.equal(d2[:, 2 * C:2 * C + K], d[:, 2 * C:2 * C + K] * 2)               # power-of-two scaling is exact


def test_gather_full_size_shift(ops, cuda):
    """Finest visit (9 channels, 864 x 864): a flow displaced by exactly one pixel to the right returns y shifted by
    one column (zeros enter at the border)."""
    side = 864
    y = torch.randn(B16, side, side, 9, device=cuda, generator=torch.Generator(device="cuda").manual_seed(6)) \
        .to(torch.bfloat16).permute(0, 3, 1, 2)
    flow = _grid(B16, side, side, cuda)
    flow[:, 0] += 2.0 / side
    got = ops.grid_sample(y, flow.permute(0, 2, 3, 1))
    assert (got[..., :-1].float() - y[..., 1:].float()).abs().max().item() < 2e-2      # fp32 coordinate rounding -> ~1e-4 weights
    assert got[..., -1].abs().max().item() < 2e-2 * y.abs().max().item() + 1e-2


# ------------------------------------------------------------------------------------------------ K3a
def test_update_resample_full_size_identity(ops, cuda):
    """864 x 864: resampling to the same size is the identity (bit for bit); the update equals the reference's
    expression evaluated by ATen on the device."""
    side = 864
    g = torch.Generator(device="cuda").manual_seed(8)
    flow = torch.randn(B16, 2, side, side, device=cuda, generator=g)
    cert = torch.randn(B16, 1, side, side, device=cuda, generator=g)
    delta = torch.randn(B16, 3, side, side, device=cuda, generator=g)
    f, c, fn, cn = ops.flow_update_resample(flow, cert, delta, 1, (side, side), (side, side))
    # (ATen divides by a host scalar as a multiplication by its reciprocal on CUDA, the kernel divides like the
    # reference's CPU path: the two differ by at most one ulp of the displacement)
    want_f = flow + 1 * torch.stack((delta[:, 0].float() / (4 * side), delta[:, 1].float() / (4 * side)), dim=1)
    assert (f - want_f).abs().max().item() < 1e-6 and torch.equal(c, cert + delta[:, 2:3])
    assert torch.equal(fn, f) and torch.equal(cn, c)


def test_epilogue_full_size_properties(ops, cuda):
    """Symmetric epilogue at 864 x 864, 8 pairs: certainty in [0, 1], warp clamped to [-1, 1], the A half of the
    warp carries the pixel grid, and out-of-range targets have certainty exactly 0."""
    side, Bp = 864, 8
    g = torch.Generator(device="cuda").manual_seed(9)
    flow = torch.randn(2 * Bp, 2, side, side, device=cuda, generator=g) * 0.8
    cert = torch.randn(2 * Bp, 1, side, side, device=cuda, generator=g) * 3
    cert16 = torch.randn(2 * Bp, 1, 40, 40, device=cuda, generator=g)
    warp, certainty = ops.match_epilogue(flow, cert, cert16, True)[:2]
    assert warp.shape == (Bp, side, 2 * side, 4) and certainty.shape == (Bp, side, 2 * side)
    assert certainty.min().item() >= 0 and certainty.max().item() <= 1
    assert warp.abs().max().item() <= 1
    grid = _grid(1, side, side, cuda)[0].permute(1, 2, 0)
    assert (warp[0, :, :side, :2] - grid).abs().max().item() < 1e-6 and (warp[0, :, side:, 2:] - grid).abs().max().item() < 1e-6
    oob = (flow[:Bp].abs() > 1).any(dim=1)
    assert certainty[:, :, :side][oob].abs().max().item() == 0


# ------------------------------------------------------------------------------------------------ K3b
def test_kde_full_size_permutation_invariance(ops, cuda):
    """N = 40000 (4 x num at num = 10000): the density of a point does not depend on the order of the points."""
    g = torch.Generator().manual_seed(10)
    centres = torch.rand(16, 4, generator=g) * 1.6 - 0.8
    x = (centres[torch.randint(0, 16, (40000,), generator=g)] + 0.06 * torch.randn(40000, 4, generator=g)).clamp(-1, 1).to(cuda)
    perm = torch.randperm(40000, generator=g).to(cuda)
    d0 = ops.kde(x, half=True).float()
    d1 = ops.kde(x[perm], half=True).float()
    # fp32 partial sums regroup under the permutation: at most one half ulp apart
    assert ((d0[perm] - d1).abs() <= d0[perm].abs().clamp(min=1.0) * 2.0 ** -10).all()
    assert d0.min().item() >= 1.0                                   # every point sees itself: exp(0) = 1


def test_sample_full_size(ops, cuda):
    """sample(warp, certainty, 10000) on one 864 x 1728 field: 10000 distinct rows of the dense warp."""
    side = 864
    g = torch.Generator(device="cuda").manual_seed(11)
    warp = torch.rand(side, 2 * side, 4, device=cuda, generator=g) * 2 - 1
    cert = torch.rand(side, 2 * side, device=cuda, generator=g)
    m, c = ops.sample(warp, cert, num=10000)
    assert m.shape == (10000, 4) and c.shape == (10000,)
    flat = warp.reshape(-1, 4)
    keys = (flat[:, 0].double() * 1e3 + flat[:, 1].double()).sort().values       # rows are distinct with probability ~1
    mk = (m[:, 0].double() * 1e3 + m[:, 1].double())
    pos = torch.searchsorted(keys, mk).clamp(max=keys.numel() - 1)
    assert torch.equal(keys[pos], mk)
    assert mk.unique().numel() == 10000


# ------------------------------------------------------------------------------------------------ (f-2)
@pytest.mark.parametrize("C,side", [(576, 216), (144, 432), (24, 864), (1144, 108)])
def test_dwconv_full_size_impulse_response(ops, cuda, C, side):
    """Depthwise 5x5 at the refiner-body shapes: a unit impulse per channel returns that channel's (flipped) taps
    exactly, zero elsewhere; image borders clip the response."""
    g = torch.Generator().manual_seed(12)
    taps = torch.randn(25, C, generator=g).to(cuda)
    bias = torch.zeros(C, device=cuda)
    x = torch.zeros(B16, side, side, C, device=cuda, dtype=torch.bfloat16).permute(0, 3, 1, 2)
    pts = [(0, 0), (side // 2, side // 3), (side - 1, side - 1), (17, side - 2)]
    for b, (py, px) in enumerate(pts):
        x[b, :, py, px] = 1.0
    out = ops.dwconv5x5_bias_act(x, taps, bias, relu=False).float()
    for b, (py, px) in enumerate(pts):
        seen = torch.zeros_like(out[b, 0], dtype=torch.bool)
        for ky in range(5):
            for kx in range(5):
                oy, ox = py - (ky - 2), px - (kx - 2)              # cross-correlation: out[o] = sum_k w[k] in[o + k - 2]
                if 0 <= oy < side and 0 <= ox < side:
                    want = taps[ky * 5 + kx].to(torch.bfloat16).float()     # output rounding of a single product
                    assert torch.equal(out[b, :, oy, ox], want)
                    seen[oy, ox] = True
        assert out[b][:, ~seen].abs().max().item() == 0
    assert out[len(pts):].abs().max().item() == 0


def test_out_conv_and_narrow_full_size_linearity(ops, cuda):
    """out_conv (576 -> 3 at 216 x 216) and the 24 -> 24 tensor-core 1x1 conv at 864 x 864: f(2x) = 2 f(x) exactly
    (bias 0), and a one-hot input channel returns that weight column."""
    g = torch.Generator(device="cuda").manual_seed(13)
    x = torch.randn(B16, 216, 216, 576, device=cuda, generator=g).to(torch.bfloat16).permute(0, 3, 1, 2)
    w3 = torch.randn(3, 576, device=cuda, generator=g)
    z3 = torch.zeros(3, device=cuda)
    a = ops.refiner_out_conv(x, w3, z3)
    assert torch.equal(ops.refiner_out_conv(x * 2, w3, z3), a * 2)
    onehot = torch.zeros(1, 4, 4, 576, device=cuda, dtype=torch.bfloat16).permute(0, 3, 1, 2)
    onehot[:, 100] = 1
    assert torch.equal(ops.refiner_out_conv(onehot, w3, z3)[0, :, 0, 0], w3[:, 100])

    y = torch.randn(B16, 864, 864, 24, device=cuda, generator=g).to(torch.bfloat16).permute(0, 3, 1, 2)
    w24 = torch.randn(24, 24, device=cuda, generator=g)
    z24 = torch.zeros(24, device=cuda)
    o = ops.pointwise_narrow(y, w24, z24)
    assert torch.equal(ops.pointwise_narrow(y * 2, w24, z24), o * 2)
    want = torch.einsum("bchw,oc->bohw", y[:1, :, :64, :64].float(), w24.to(torch.bfloat16).float())
    assert (o[:1, :, :64, :64].float() - want).abs().max().item() < 2e-2 * want.abs().max().item()


def test_maxpool_layernorm_full_size(ops, cuda):
    """VGG stage pooling at 864 x 864 x 64 is bit-identical to ATen; LayerNorm output rows have zero mean / unit
    variance for gamma = 1, beta = 0 at the DINOv2 token shape (16, 1601, 1024)."""
    g = torch.Generator(device="cuda").manual_seed(14)
    x = torch.randn(B16, 864, 864, 64, device=cuda, generator=g).to(torch.bfloat16).permute(0, 3, 1, 2)
    assert torch.equal(ops.maxpool2x2(x), torch.nn.functional.max_pool2d(x, 2, 2))
    t = (torch.randn(B16, 1601, 1024, device=cuda, generator=g) * 3 + 1).to(torch.bfloat16)
    n = ops.layer_norm(t, torch.ones(1024, device=cuda, dtype=torch.bfloat16), torch.zeros(1024, device=cuda, dtype=torch.bfloat16), 1e-6).float()
    assert n.mean(-1).abs().max().item() < 5e-3 and (n.var(-1, unbiased=False) - 1).abs().max().item() < 1e-2


# ------------------------------------------------------------------------------------------------ whole path
def test_match_full_size_swap_symmetry(cuda):
    """roma_outdoor 560 -> 864 in bf16, 2 pairs: matching (A, B) and (B, A) exchange the two halves of the symmetric
    warp (the decoder treats the two orderings as the two halves of one batch), outputs are finite and in range."""
    import roma_b200
    weights, dinov2 = roma_b200.random_init_weights(seed=0)
    model = roma_b200.roma_outdoor(cuda, weights, dinov2, amp_dtype=torch.bfloat16)
    g = torch.Generator().manual_seed(15)
    A = (torch.rand(2, 3, 864, 864, generator=g) * 255).to(torch.uint8)
    Bm = (torch.rand(2, 3, 864, 864, generator=g) * 255).to(torch.uint8)
    w_ab, c_ab = model.match(A, Bm, batched=True, device=cuda)
    w_ba, c_ba = model.match(Bm, A, batched=True, device=cuda)
    assert w_ab.shape == (2, 864, 1728, 4) and torch.isfinite(w_ab).all() and torch.isfinite(c_ab).all()
    assert w_ab.abs().max().item() <= 1 and 0 <= c_ab.min().item() and c_ab.max().item() <= 1
    # A->B half of (A,B) == the B->A half of (B,A) with source / target coordinates exchanged
    assert (w_ab[:, :, :864, 2:] - w_ba[:, :, 864:, :2]).abs().max().item() < 2e-2      # bf16 bodies, batch-order noise
    assert (c_ab[:, :, :864] - c_ba[:, :, 864:]).abs().max().item() < 5e-2
    m, c = model.sample(w_ab[0], c_ab[0], num=10000)
    assert m.shape == (10000, 4) and c.shape == (10000,)
    # sample_batch == sample per pair with the same generators (two streams inside, same RNG consumption)
    g1 = [torch.Generator(device=cuda).manual_seed(100 + i) for i in range(2)]
    g2 = [torch.Generator(device=cuda).manual_seed(100 + i) for i in range(2)]
    ms, cs = model.sample_batch(w_ab, c_ab, num=5000, generators=g1)
    for i in range(2):
        mi, ci = model.sample(w_ab[i], c_ab[i], num=5000, generator=g2[i])
        assert torch.equal(ms[i], mi) and torch.equal(cs[i], ci)


# ------------------------------------------------------------------------------------------------ config 4 (672 -> 1344)
def test_gp_config4_n2304_vs_fp64_oracle(ops, cuda):
    """BASELINE config 4: the GP at 48 x 48 = 2304 tokens (18 diagonal blocks of the blocked Cholesky) against the
    fp64 oracle, both schedules, fp32 and bf16 features."""
    from roma_b200 import _lib
    lib = _lib.load()
    g = torch.Generator().manual_seed(4)
    base = O.bilinear_resize(torch.randn(1, 512, 6, 6, generator=g), (48, 48))
    x32 = base + 0.5 * torch.randn(1, 512, 48, 48, generator=g)
    y32 = base * 0.7 + 0.5 * torch.randn(1, 512, 48, 48, generator=g)
    pw = torch.randn(512, 2, generator=g) * 0.5
    pb = torch.randn(512, generator=g) * 0.5
    for dt in (torch.bfloat16, torch.float32):
        x, y = x32.to(dt), y32.to(dt)
        want = O.gp_posterior(x.float(), y.float(), pw, pb, dtype=torch.float64)
        try:
            for lanes in (1, 0):
                lib.roma_gp_set_lanes(lanes)
                got = ops.gp_posterior(x.repeat(3, 1, 1, 1).to(cuda), y.repeat(3, 1, 1, 1).to(cuda), pw.to(cuda), pb.to(cuda))
                assert got.shape == (3, 512, 48, 48)
                assert (got[0].float().cpu() - want[0].float()).abs().max().item() < 1e-3, (dt, lanes)
                assert torch.equal(got[1], got[0]) and torch.equal(got[2], got[0])
        finally:
            lib.roma_gp_set_lanes(1)


@pytest.mark.parametrize("coarse,up", [((672, 672), (1344, 1344)), ((560, 560), (864, 1152))])
def test_match_config4_and_nonsquare_swap_symmetry(cuda, coarse, up):
    """BASELINE config 4 (672 -> 1344: local correlation at 168^2 / 336^2, gather-only at 672^2 / 1344^2) and the
    reference demo's non-square upsample resolution (demo/demo_match.py:28: 864 x 1152), one pair in bf16:
    shapes, ranges, the grid half of the warp, and (A, B) / (B, A) swap symmetry."""
    import roma_b200
    weights, dinov2 = roma_b200.random_init_weights(seed=0)
    model = roma_b200.roma_outdoor(cuda, weights, dinov2, coarse_res=coarse, upsample_res=up, amp_dtype=torch.bfloat16)
    H, W = up
    g = torch.Generator().manual_seed(16)
    A = (torch.rand(1, 3, H, W, generator=g) * 255).to(torch.uint8)
    Bm = (torch.rand(1, 3, H, W, generator=g) * 255).to(torch.uint8)
    w_ab, c_ab = model.match(A, Bm, batched=True, device=cuda)
    w_ba, c_ba = model.match(Bm, A, batched=True, device=cuda)
    assert w_ab.shape == (1, H, 2 * W, 4) and c_ab.shape == (1, H, 2 * W)
    assert torch.isfinite(w_ab).all() and torch.isfinite(c_ab).all()
    assert w_ab.abs().max().item() <= 1 and 0 <= c_ab.min().item() and c_ab.max().item() <= 1
    grid = _grid(1, H, W, cuda).permute(0, 2, 3, 1)
    assert (w_ab[:, :, :W, :2] - grid).abs().max().item() < 1e-6 and (w_ab[:, :, W:, 2:] - grid).abs().max().item() < 1e-6
    assert (w_ab[:, :, :W, 2:] - w_ba[:, :, W:, :2]).abs().max().item() < 2e-2
    assert (c_ab[:, :, :W] - c_ba[:, :, W:]).abs().max().item() < 5e-2
    m, c = model.sample(w_ab[0], c_ab[0], num=5000)
    assert m.shape == (5000, 4) and c.shape == (5000,)
