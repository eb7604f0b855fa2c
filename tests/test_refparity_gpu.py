"""GPU: parity AT THE BASELINE SIZES against the reference's own formulation executed on the same B200.

The comparand is the UNMODIFIED reference from baseline/_ref (it travels with the gpurun snapshot; oracle/reference.py) —
`local_correlation`, `F.grid_sample`, `GP.forward`, `F.interpolate`, the `match` epilogue, `roma_outdoor` itself — run on
CUDA in fp32 (TF32 off).  Where the reference is not on the machine the oracle's ATen mode (`O.set_aten(True)`: the same
formulation, restated) stands in, so the tests never silently skip.

What is checked that the property tests of test_fullsize_gpu.py cannot: bilinear weights and tap order on INCOHERENT
flows at C = 512, r = 3, 108 x 108, every one of the nine refiner visits of a 560 -> 864 match at decoder batch 16, the
560 -> 108 / 40 -> 864 resamplings, the 864 x 864 epilogue, the GP at b = 16, N = 1600, and the whole match at 560 -> 864
(fp32 bodies: north_star's tolerances; bf16 / fp16 bodies: deviation measured, printed and bounded).
Every product call goes through the C ABI (roma_b200.ops / the roma_b200 model)."""
import math

import pytest
import torch
import torch.nn.functional as F

from oracle import reference as R
from oracle import roma_oracle as O

pytestmark = pytest.mark.gpu

B16 = 16
# (scale, C, side, r, E) for the nine refiner visits of one 560 -> 864 match
VISITS = [("s16", 512, 40, 7, 128), ("s8", 512, 70, 3, 64), ("s4", 256, 140, 2, 32), ("s2", 64, 280, 0, 16),
          ("s1", 9, 560, 0, 6), ("s8u", 512, 108, 3, 64), ("s4u", 256, 216, 2, 32), ("s2u", 64, 432, 0, 16),
          ("s1u", 9, 864, 0, 6)]
EPS = {torch.float32: 2.0 ** -23, torch.bfloat16: 2.0 ** -8, torch.float16: 2.0 ** -11}


@pytest.fixture(scope="module")
def ops(cuda):
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    from roma_b200 import ops as _ops
    return _ops


@pytest.fixture(scope="module")
def ref():
    """(local_correlation, what) — the reference's function when baseline/_ref is here, else the oracle's ATen mode."""
    if R.available():
        R.import_reference()
        from romatch.utils.local_correlation import local_correlation
        return local_correlation, "baseline/_ref"

    def lc(f0, f1, local_radius, flow=None, sample_mode="bilinear"):
        O.set_aten(True)
        try:
            return O.local_correlation(f0, f1, local_radius, flow=flow)
        finally:
            O.set_aten(False)
    return lc, "oracle (ATen mode; reference not on this machine)"


def _grid(B, h, w, dev):
    ys = torch.linspace(-1 + 1 / h, 1 - 1 / h, h, device=dev)
    xs = torch.linspace(-1 + 1 / w, 1 - 1 / w, w, device=dev)
    return torch.stack((xs[None, :].expand(h, w), ys[:, None].expand(h, w)))[None].expand(B, 2, h, w).contiguous()


def _flow(kind, B, h, w, dev, g):
    """coherent = pixel grid + upsampled N(0, 0.05) low-res displacement; incoherent = U(-1.1, 1.1)."""
    if kind == "incoherent":
        return torch.rand(B, 2, h, w, device=dev, generator=g) * 2.2 - 1.1
    low = torch.randn(B, 2, max(h // 8, 2), max(w // 8, 2), device=dev, generator=g) * 0.05
    return _grid(B, h, w, dev) + F.interpolate(low, size=(h, w), mode="bilinear", align_corners=False)


def _bound(want, dtype, atol):
    """|got - want| allowed: rounding of the output to `dtype` (half an ulp, 0.51 for the fp32 arithmetic underneath) + atol."""
    return 0.51 * EPS[dtype] * 2 * want.abs() + atol


@pytest.mark.parametrize("kind", ["incoherent", "coherent"])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32], ids=["bf16", "fp32"])
@pytest.mark.parametrize("visit", VISITS, ids=[v[0] for v in VISITS])
def test_refiner_inputs_teacher_forced_vs_reference(ops, ref, cuda, visit, dtype, kind):
    """matcher.py:105-132 at decoder batch 16: d = [x | grid_sample(y, flow) | disp_emb | local_correlation] from ONE launch
    against the reference's four ops on the fp32 values of the same features (so the comparison isolates the kernel:
    16-bit outputs must be the correctly rounded fp32 results up to the fp32 summation order)."""
    local_correlation, what = ref
    name, C, side, r, E = visit
    B = B16 if dtype != torch.float32 or side <= 216 else 4          # fp32 at 280..864: 4 elements keep the test short
    g = torch.Generator(device="cuda").manual_seed(2 * VISITS.index(visit) + (kind == "coherent"))
    x = torch.randn(B, side, side, C, device=cuda, generator=g).to(dtype).permute(0, 3, 1, 2)
    y = torch.randn(B, side, side, C, device=cuda, generator=g).to(dtype).permute(0, 3, 1, 2)
    flow = _flow(kind, B, side, side, cuda, g)
    dw = torch.randn(E, 2, 1, 1, device=cuda, generator=g)
    db = torch.randn(E, device=cuda, generator=g)
    scale_factor = 1.0 if not name.endswith("u") else math.sqrt(864 * 864 / (560 * 560))
    d = ops.refiner_inputs(x, y, flow, dw, db, 40 / 32 * scale_factor, r or None, pad_to=8)
    K = (2 * r + 1) ** 2 if r else 0
    assert d.shape == (B, (2 * C + E + K + 7) // 8 * 8, side, side) and d.dtype == dtype
    xf, yf = x.float(), y.float()
    with torch.no_grad():
        xhat = F.grid_sample(yf, flow.permute(0, 2, 3, 1), align_corners=False, mode="bilinear")
        emb = F.conv2d(40 / 32 * scale_factor * (flow - _grid(B, side, side, cuda)), dw, db)
    assert torch.equal(d[:, :C], x)
    e_xhat = (d[:, C:2 * C].float() - xhat).abs()
    assert bool((e_xhat <= _bound(xhat, dtype, 2e-5)).all()), f"x_hat max err {e_xhat.max().item():.3e}"
    e_emb = (d[:, 2 * C:2 * C + E].float() - emb).abs()
    assert bool((e_emb <= _bound(emb, dtype, 2e-5)).all()), f"emb max err {e_emb.max().item():.3e}"
    msg = f"{name} {kind} {str(dtype)[6:]} B={B}: x_hat {e_xhat.max().item():.2e} emb {e_emb.max().item():.2e}"
    if r:
        e_corr = torch.zeros((), device=cuda)
        e_own = torch.zeros((), device=cuda)
        for b0 in range(0, B, 4):                                    # the reference materialises (C, h, w K) per element
            sl = slice(b0, b0 + 4)
            with torch.no_grad():
                want = local_correlation(xf[sl].contiguous(), yf[sl].contiguous(), local_radius=r, flow=flow[sl], sample_mode="bilinear")
            got = d[sl, 2 * C + E:2 * C + E + K].float()
            err = (got - want).abs()
            # atol: the reference un-normalises every window tap separately (flow + k * 2 / w, rounded in fp32, then * w / 2), so
            # its 4 bilinear weights differ from tap to tap by ~1e-5 and a tap within that of a pixel boundary switches
            # neighbours; the kernel uses the centre's weights for all taps.  Measured max over 18 M values: 1.5e-4
            assert bool((err <= _bound(want, dtype, 3e-4)).all()), f"corr max err {err.max().item():.3e} at {name} {kind}"
            e_corr = torch.maximum(e_corr, err.max())
            if dtype != torch.float32 and b0 == 0:                   # for the record: the reference's own 16-bit chain vs its fp32 one
                # as ConvRefiner.forward runs it: under autocast, where grid_sample is on the fp32 list (16-bit features are
                # up-cast for the sampling, the products and the sum; the result is stored in the feature dtype)
                with torch.no_grad(), torch.autocast("cuda", dtype=dtype):
                    own = local_correlation(x[sl].contiguous(), y[sl].contiguous(), local_radius=r, flow=flow[sl], sample_mode="bilinear")
                e_own = (own.float() - want).abs().max()
        msg += f" corr {e_corr.item():.2e}" + (f" (reference's own {str(dtype)[6:]} chain vs its fp32: {e_own.item():.2e})" if dtype != torch.float32 else "")
    assert d[:, 2 * C + E + K:].abs().max().item() == 0 if d.shape[1] > 2 * C + E + K else True
    print(msg + f"   [{what}]")


@pytest.mark.parametrize("src,dst", [(560, 108), (40, 864), (108, 216), (40, 70)])
def test_resample_vs_aten_at_baseline_sizes(ops, cuda, src, dst):
    """matcher.py:349-360 / :407-417 / :657: F.interpolate(bilinear, align_corners=False) of flow and certainty, B = 16."""
    g = torch.Generator(device="cuda").manual_seed(src + dst)
    flow = torch.randn(B16, 2, src, src, device=cuda, generator=g)
    cert = torch.randn(B16, 1, src, src, device=cuda, generator=g) * 3
    _, _, fn, cn = ops.flow_update_resample(flow, cert, None, 1, (864, 864), (dst, dst), want_current=False)
    wf = F.interpolate(flow, size=(dst, dst), mode="bilinear", align_corners=False)
    wc = F.interpolate(cert, size=(dst, dst), mode="bilinear", align_corners=False)
    ef, ec = (fn - wf).abs().max().item(), (cn - wc).abs().max().item()
    print(f"resample {src} -> {dst}: flow {ef:.2e} certainty {ec:.2e}")
    assert ef < 1e-5 and ec < 3e-5


def test_update_vs_reference_expression_at_864(ops, cuda):
    """matcher.py:397-402 at the finest visit, B = 16: the kernel follows the reference's CPU rounding (true division);
    ATen CUDA multiplies by the reciprocal — at most one ulp of the displacement apart."""
    g = torch.Generator(device="cuda").manual_seed(11)
    side, ins = 864, 1
    flow = torch.randn(B16, 2, side, side, device=cuda, generator=g)
    cert = torch.randn(B16, 1, side, side, device=cuda, generator=g)
    delta = torch.randn(B16, 3, side, side, device=cuda, generator=g)
    f, c, _, _ = ops.flow_update_resample(flow, cert, delta, ins, (side, side), None)
    disp = ins * torch.stack((delta[:, 0].float() / (4 * side), delta[:, 1].float() / (4 * side)), dim=1)
    assert (f - (flow + disp)).abs().max().item() < 2e-7 * 4 and torch.equal(c, cert + delta[:, 2:3])


def test_epilogue_vs_reference_expression_at_864(ops, cuda):
    """matcher.py:656-662, 684-718 at 864 x 864, 8 pairs (the reference's own lines, run by ATen on the device)."""
    side, Bp = 864, 8
    g = torch.Generator(device="cuda").manual_seed(12)
    flow = torch.randn(2 * Bp, 2, side, side, device=cuda, generator=g) * 0.7
    cert = torch.randn(2 * Bp, 1, side, side, device=cuda, generator=g) * 3
    cert16 = torch.randn(2 * Bp, 1, 40, 40, device=cuda, generator=g) * 2
    warp, prob, logits = ops.match_epilogue(flow, cert, cert16, True, return_logits=True)
    low = F.interpolate(cert16, size=(side, side), align_corners=False, mode="bilinear")
    low = 0.5 * low * (low < 0)
    want_logits = cert - low
    want = want_logits.sigmoid()
    fl = flow.permute(0, 2, 3, 1)
    want = torch.where((fl.abs() > 1).any(-1)[:, None], torch.zeros_like(want), want)
    fl = fl.clamp(-1, 1)
    grid = _grid(Bp, side, side, cuda).permute(0, 2, 3, 1)
    want_warp = torch.cat((torch.cat((grid, fl[:Bp]), -1), torch.cat((fl[Bp:], grid), -1)), dim=2)
    want_prob = torch.cat((want[:Bp, 0], want[Bp:, 0]), dim=2)
    want_log = torch.cat((want_logits[:Bp, 0], want_logits[Bp:, 0]), dim=2)
    ew, ep, el = (warp - want_warp).abs().max().item(), (prob - want_prob).abs().max().item(), (logits - want_log).abs().max().item()
    print(f"epilogue 864^2 x 8 pairs: warp {ew:.2e} certainty {ep:.2e} logits {el:.2e}")
    assert ew < 1e-6 and ep < 1e-5 and el < 3e-5


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32], ids=["bf16", "fp32"])
def test_gp_b16_n1600_vs_reference_gp_forward(ops, cuda, dtype):
    """matcher.py:249-273 at b = 16, N = 1600, C = D = 512: the reference's GP.forward (fp32 LU inverse, TF32 off) on CUDA."""
    g = torch.Generator(device="cuda").manual_seed(13)
    base = F.interpolate(torch.randn(B16, 512, 6, 6, device=cuda, generator=g), size=(40, 40), mode="bilinear")
    x = (base + 0.5 * torch.randn(B16, 512, 40, 40, device=cuda, generator=g)).to(dtype)
    y = (base.roll(1, 0) * 0.7 + 0.5 * torch.randn(B16, 512, 40, 40, device=cuda, generator=g)).to(dtype)
    if R.available():
        R.import_reference()
        from romatch.models.matcher import GP, CosKernel
        gp = GP(CosKernel, T=0.2, learn_temperature=False, only_attention=False, gp_dim=512, basis="fourier", no_cov=True).to(cuda).eval()
        pw, pb = gp.pos_conv.weight.detach(), gp.pos_conv.bias.detach()
        with torch.no_grad():
            want = gp(x, y)
        what = "baseline/_ref GP.forward"
    else:
        pw = torch.randn(512, 2, 1, 1, device=cuda, generator=g) * 0.5
        pb = torch.randn(512, device=cuda, generator=g) * 0.5
        want = torch.cat([O.gp_posterior(x[i:i + 1].float(), y[i:i + 1].float(), pw.reshape(512, 2), pb) for i in range(B16)])
        what = "oracle gp_posterior on CUDA"
    got = ops.gp_posterior(x, y, pw, pb)
    err = (got - want).abs().max().item()
    # fp64 truth for one element: how far the reference's own fp32 path is from it (the yardstick for `err`)
    t64 = O.gp_posterior(x[:1].float().cpu(), y[:1].float().cpu(), pw.reshape(512, 2).cpu(), pb.cpu(), dtype=torch.float64)
    e_ref = (want[:1].cpu().double() - t64).abs().max().item()
    e_own = (got[:1].cpu().double() - t64).abs().max().item()
    print(f"GP b=16 N=1600 {str(dtype)[6:]}: vs {what} {err:.2e}; vs fp64: reference {e_ref:.2e}, kernel {e_own:.2e}")
    assert got.shape == want.shape == (B16, 512, 40, 40)
    assert err < 1e-3 and e_own < 1e-3                                          # north_star's budget for this stage
