"""Parity of every CUDA entry point (through the C ABI) against the CPU oracle and the reference goldens.

Tolerances: north_star asks for warp <= 1e-3 and certainty logits <= 1e-2 end to end; per kernel we hold
fp32 paths to ~1e-5..1e-4 (accumulation-order noise) and 16-bit feature paths to the rounding of their
output dtype.  Integer outputs (indices) are compared exactly.
"""
import math

import pytest
import torch

from oracle import roma_oracle as O

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ops(cuda):
    from roma_b200 import _lib, ops as _ops
    lib = _lib.load()
    assert lib.roma_check_device(0) == 0, lib.roma_last_error_string()
    return _ops


def err(a, b):
    return (a.float().cpu() - b.float().cpu()).abs().max().item()


def make_flow(g, B, h, w, mode):
    grid = O.pixel_grid(B, h, w)
    if mode == "coherent":
        low = torch.randn(B, 2, max(h // 8, 2), max(w // 8, 2), generator=g) * 0.05
        return (grid + O.bilinear_resize(low, (h, w))).contiguous()
    return torch.rand(B, 2, h, w, generator=g) * 2.2 - 1.1      # incoherent, includes out-of-range taps


# ---------------------------------------------------------------------------------------------- K2
@pytest.mark.parametrize("tag", ["r2", "r3", "r7"])
def test_local_corr_golden(ops, cuda, golden, tag):
    """Kernel vs the REFERENCE's own outputs (tests/golden), fp32."""
    g = golden(f"local_corr_{tag}")
    # golden channel counts (16, 24, 8) are multiples of 4 -> vector path in fp32
    corr = ops.local_correlation(g["f0"].to(cuda), g["f1"].to(cuda), g["r"], flow=g["flow"].to(cuda))
    assert err(corr, g["corr"]) < 1e-4
    xhat = ops.grid_sample(g["f1"].to(cuda), g["flow"].permute(0, 2, 3, 1).to(cuda))
    assert err(xhat, g["xhat"]) < 1e-5


def test_local_corr_aligned_golden(ops, cuda, golden):
    g = golden("local_corr_noflow")
    assert err(ops.local_correlation(g["f0"].to(cuda), g["f1"].to(cuda), g["r"]), g["corr"]) < 1e-4


@pytest.mark.parametrize("tag", ["corr", "nocorr"])
def test_refiner_inputs_golden(ops, cuda, golden, tag):
    g = golden(f"refiner_inputs_{tag}")
    d = ops.refiner_inputs(g["x"].to(cuda), g["y"].to(cuda), g["flow"].to(cuda), g["disp_w"].to(cuda), g["disp_b"].to(cuda),
                           40 / 32 * g["scale_factor"], g["r"] or None)
    assert d.shape == g["d"].shape
    assert err(d, g["d"]) < 1e-4


@pytest.mark.parametrize("dtype,tol", [(torch.float32, 2e-4), (torch.float16, 6e-3), (torch.bfloat16, 4e-2)])
@pytest.mark.parametrize("C,E,r,h,w", [(512, 128, 7, 20, 20), (512, 64, 3, 35, 33), (256, 32, 2, 40, 44),
                                        (64, 16, 0, 48, 40), (9, 6, 0, 50, 46), (24, 0, 0, 30, 40)])
@pytest.mark.parametrize("flow_mode", ["coherent", "incoherent"])
def test_refiner_inputs_vs_oracle(ops, cuda, dtype, tol, C, E, r, h, w, flow_mode):
    """The five roma_outdoor refiner configurations (channel counts of roma_models.py:40-114) + a Tiny gather,
    at reduced spatial size, in the three feature dtypes, for coherent and incoherent flows."""
    g = torch.Generator().manual_seed(7)
    B = 2
    x = torch.randn(B, C, h, w, generator=g).to(dtype)
    y = torch.randn(B, C, h, w, generator=g).to(dtype)
    flow = make_flow(g, B, h, w, flow_mode)
    dw = torch.randn(E, 2, generator=g) if E else None
    db = torch.randn(E, generator=g) if E else None
    want = O.refiner_inputs(x.float(), y.float(), flow, dw, db, 1.9, r or None)      # oracle on the SAME rounded inputs
    got = ops.refiner_inputs(x.to(cuda), y.to(cuda), flow.to(cuda), None if dw is None else dw.to(cuda),
                             None if db is None else db.to(cuda), 1.9, r or None)
    assert got.dtype == dtype and got.shape == want.shape
    assert got.is_contiguous(memory_format=torch.channels_last)
    scale = max(1.0, want.abs().max().item())
    assert err(got, want) < tol * scale


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("C,r,h,w", [(512, 7, 12, 14), (512, 3, 23, 19), (256, 2, 31, 27), (128, 1, 17, 20), (64, 4, 16, 16)])
@pytest.mark.parametrize("flow_mode", ["coherent", "incoherent"])
def test_local_corr_tensor_core_vs_simt_and_oracle(ops, cuda, dtype, C, r, h, w, flow_mode):
    """Aligned 16-bit local correlation: the tensor-core dot products (mma.sync m16n8k16, default) against the SIMT
    FMA kernel and the fp32 oracle on the same rounded features — correlation channels agree to the rounding of the
    16-bit output; the copied / gathered segments (x, x_hat) are bit-identical between the two kernels."""
    from roma_b200 import _lib
    lib = _lib.load()
    g = torch.Generator().manual_seed(31)
    B = 3
    x = torch.randn(B, C, h, w, generator=g).to(dtype)
    y = (0.6 * x.roll(1, 3).float() + 0.8 * torch.randn(B, C, h, w, generator=g)).to(dtype)
    flow = make_flow(g, B, h, w, flow_mode)
    want = O.refiner_inputs(x.float(), y.float(), flow, None, None, 1.0, r)
    try:
        assert lib.roma_local_corr_set_path(1) == 0
        simt = ops.refiner_inputs(x.to(cuda), y.to(cuda), flow.to(cuda), None, None, 1.0, r, pad_to=8)
        assert lib.roma_local_corr_set_path(0) == 0
        tc = ops.refiner_inputs(x.to(cuda), y.to(cuda), flow.to(cuda), None, None, 1.0, r, pad_to=8)
    finally:
        lib.roma_local_corr_set_path(0)
    K = (2 * r + 1) ** 2
    assert tc.shape == simt.shape and tc.shape[1] == (want.shape[1] + 7) // 8 * 8     # padded pitch: the aligned kernels
    assert tc[:, want.shape[1]:].abs().max().item() == 0 if tc.shape[1] > want.shape[1] else True
    assert torch.equal(tc[:, : 2 * C], simt[:, : 2 * C])
    ulp = 2.0 ** (-10 if dtype == torch.float16 else -7)
    scale = max(1.0, want[:, 2 * C: 2 * C + K].abs().max().item())
    assert err(tc[:, 2 * C: 2 * C + K], want[:, 2 * C: 2 * C + K]) < 1.5 * ulp * scale
    assert err(tc[:, 2 * C: 2 * C + K], simt[:, 2 * C: 2 * C + K]) < 1.5 * ulp * scale


def test_refiner_inputs_strided_flow_and_nchw_features(ops, cuda):
    """Permuted flow views and NCHW-contiguous features are accepted (SURVEY §8a 'strided inputs')."""
    g = torch.Generator().manual_seed(3)
    x = torch.randn(2, 64, 12, 10, generator=g); y = torch.randn(2, 64, 12, 10, generator=g)
    flow_nhwc = make_flow(g, 2, 12, 10, "incoherent").permute(0, 2, 3, 1).contiguous()
    flow = flow_nhwc.permute(0, 3, 1, 2)   # (B,2,h,w) view over interleaved storage
    want = O.refiner_inputs(x, y, flow, None, None, 1.0, 2)
    got = ops.refiner_inputs(x.to(cuda), y.to(cuda), flow_nhwc.to(cuda).permute(0, 3, 1, 2), None, None, 1.0, 2)
    assert err(got, want) < 1e-4


def test_refiner_gather_rejects_cpu_tensors(ops):
    with pytest.raises(RuntimeError):
        ops.refiner_inputs(torch.zeros(1, 8, 4, 4), torch.zeros(1, 8, 4, 4), torch.zeros(1, 2, 4, 4))


# --------------------------------------------------------------------------------------------- K3a
def test_update_resample_golden(ops, cuda, golden):
    g = golden("update_resample")
    for k in ("up", "odd", "down"):
        size = tuple(int(v) for v in g[f"size_{k}"])
        fo, co, fn, cn = ops.flow_update_resample(g["flow"].to(cuda), g["cert"].to(cuda), g["delta"].to(cuda),
                                                  g["ins"], (g["Hf"], g["Wf"]), size)
        assert torch.equal(fo.cpu(), g["flow_u"]) and torch.equal(co.cpu(), g["cert_u"])   # update is bit-exact
        assert err(fn, g[f"flow_{k}"]) < 1e-5 and err(cn, g[f"cert_{k}"]) < 1e-5


@pytest.mark.parametrize("h,w,h2,w2", [(40, 40, 70, 70), (70, 70, 140, 140), (560, 560, 108, 108), (36, 52, 63, 91)])
def test_update_resample_vs_oracle(ops, cuda, h, w, h2, w2):
    g = torch.Generator().manual_seed(11)
    B = 2
    flow = make_flow(g, B, h, w, "coherent"); cert = torch.randn(B, 1, h, w, generator=g) * 2
    delta = (torch.randn(B, h, w, 3, generator=g) * 4).permute(0, 3, 1, 2)     # channels-last delta like out_conv gives
    fu, cu = O.flow_update(flow, cert, delta, 8, 4 * h, 4 * w)
    fo, co, fn, cn = ops.flow_update_resample(flow.to(cuda), cert.to(cuda), delta.to(cuda), 8, (4 * h, 4 * w), (h2, w2))
    assert torch.equal(fo.cpu(), fu) and torch.equal(co.cpu(), cu)
    assert err(fn, O.bilinear_resize(fu, (h2, w2))) < 1e-5
    assert err(cn, O.bilinear_resize(cu, (h2, w2))) < 5e-5   # |values| up to ~15: a few fp32 ulps of the lerp weights
    # pure resample, no certainty (entry of the coarse pass / upsample pass)
    _, _, fn2, cn2 = ops.flow_update_resample(flow.to(cuda), None, None, 8, (h, w), (h2, w2), want_current=False)
    assert err(fn2, O.bilinear_resize(flow, (h2, w2))) < 1e-5 and cn2.abs().max().item() == 0


@pytest.mark.parametrize("tag", ["sym", "asym"])
def test_epilogue_golden(ops, cuda, golden, tag):
    g = golden(f"epilogue_{tag}")
    warp, cert = ops.match_epilogue(g["flow"].to(cuda), g["cert"].to(cuda),
                                    g["cert16"].to(cuda) if g["attenuate"] else None, bool(g["symmetric"]))
    assert err(warp[0], g["warp"]) < 1e-6 and err(cert[0], g["certainty"]) < 1e-6


def test_epilogue_vs_oracle_batched(ops, cuda):
    g = torch.Generator().manual_seed(5)
    B, H, W = 3, 96, 80
    flow = make_flow(g, 2 * B, H, W, "coherent") * 1.05        # some coordinates beyond +-1
    cert = torch.randn(2 * B, 1, H, W, generator=g) * 3
    c16 = torch.randn(2 * B, 1, 7, 5, generator=g) * 2
    w0, c0, l0 = O.match_epilogue(flow, cert, c16, True)
    w1, c1, l1 = ops.match_epilogue(flow.to(cuda), cert.to(cuda), c16.to(cuda), True, return_logits=True)
    assert w1.shape == (B, H, 2 * W, 4)
    assert err(w1, w0) < 1e-6 and err(c1, c0) < 1e-6 and err(l1, l0) < 1e-5
    assert (c1 == 0).sum().item() == (c0 == 0).sum().item() > 0


# ---------------------------------------------------------------------------------------------- K1
@pytest.mark.parametrize("tag", ["sq", "rect"])
def test_gp_golden(ops, cuda, golden, tag):
    g = golden(f"gp_{tag}")
    mu = ops.gp_posterior(g["x"].to(cuda), g["y"].to(cuda), g["pos_w"].to(cuda), g["pos_b"].to(cuda))
    assert mu.shape == g["mu"].shape
    assert err(mu, g["mu"]) < 1e-4


@pytest.mark.parametrize("h1,w1,h2,w2,C,D", [(20, 20, 20, 20, 512, 512), (12, 17, 15, 11, 128, 96)])
def test_gp_vs_oracle_fp64(ops, cuda, h1, w1, h2, w2, C, D):
    """Error budget: |mu - mu_fp64| <= 1e-3 (GP error feeds the warp with gain ~0.55)."""
    g = torch.Generator().manual_seed(21)
    b = 2
    base = O.bilinear_resize(torch.randn(b, C, 4, 4, generator=g), (h1, w1))
    x = base + 0.3 * torch.randn(b, C, h1, w1, generator=g)
    y = O.bilinear_resize(base, (h2, w2)) + 0.3 * torch.randn(b, C, h2, w2, generator=g)
    pw = torch.randn(D, 2, generator=g) * 0.5; pb = torch.randn(D, generator=g) * 0.5
    want = O.gp_posterior(x, y, pw, pb, dtype=torch.float64)
    ref32 = O.gp_posterior(x, y, pw, pb)
    got = ops.gp_posterior(x.to(cuda), y.to(cuda), pw.to(cuda), pb.to(cuda))
    e_kernel, e_ref = err(got, want), err(ref32, want)
    assert e_kernel < 1e-3, (e_kernel, e_ref)
    assert e_kernel < max(10 * e_ref, 2e-4), (e_kernel, e_ref)      # as good as the reference's own fp32 path


@pytest.mark.parametrize("path", [2, 3])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.float16])
@pytest.mark.parametrize("h1,w1,h2,w2", [(16, 16, 16, 16), (13, 21, 19, 15), (40, 40, 40, 40)])
def test_gp_tcgen05_path(ops, cuda, dtype, h1, w1, h2, w2, path):
    """The tcgen05/TMEM pipeline (forced) vs the fp64 oracle and vs the fp32 SIMT path, on the SAME (already
    rounded) features: 16-bit inputs are exact in the tensor-core Gram, fp32 inputs go through split-bf16."""
    from roma_b200 import _lib
    lib = _lib.load()
    g = torch.Generator().manual_seed(33)
    b, C, D = 2, 512, 512
    base = O.bilinear_resize(torch.randn(b, C, 5, 5, generator=g), (h1, w1))
    x = (base + 0.4 * torch.randn(b, C, h1, w1, generator=g)).to(dtype)
    y = (O.bilinear_resize(base, (h2, w2)) + 0.4 * torch.randn(b, C, h2, w2, generator=g)).to(dtype)
    pw = torch.randn(D, 2, generator=g) * 0.5; pb = torch.randn(D, generator=g) * 0.5
    want = O.gp_posterior(x.float(), y.float(), pw, pb, dtype=torch.float64)
    ref32 = O.gp_posterior(x.float(), y.float(), pw, pb)
    try:
        assert lib.roma_gp_set_path(path) == 0        # 2: 3-plane solve GEMMs (the default), 3: 2-plane (opt-in)
        got_tc = ops.gp_posterior(x.to(cuda), y.to(cuda), pw.to(cuda), pb.to(cuda))
        assert lib.roma_gp_set_path(1) == 0
        got_simt = ops.gp_posterior(x.to(cuda), y.to(cuda), pw.to(cuda), pb.to(cuda))
    finally:
        lib.roma_gp_set_path(0)
    e_tc, e_simt, e_ref = err(got_tc, want), err(got_simt, want), err(ref32, want)
    assert e_tc < max(10 * e_ref, 2e-4), (e_tc, e_simt, e_ref)
    assert err(got_tc, got_simt) < 5e-4


@pytest.mark.parametrize("h1,w1,h2,w2,b", [(40, 40, 40, 40, 4), (13, 21, 19, 15, 2), (30, 30, 24, 16, 3)])
def test_gp_two_lane_schedule(ops, cuda, h1, w1, h2, w2, b):
    """The two-lane schedule (chain diag -> panel -> syrk on the library's helper stream, bulk updates one step
    behind on the caller's stream) against the single-stream left-looking schedule and the fp64 oracle; repeated
    calls on a non-default stream are bit-reproducible (no race between the lanes or between calls)."""
    from roma_b200 import _lib
    lib = _lib.load()
    g = torch.Generator().manual_seed(77)
    C, D = 512, 512
    base = O.bilinear_resize(torch.randn(b, C, 5, 5, generator=g), (h1, w1))
    x = (base + 0.4 * torch.randn(b, C, h1, w1, generator=g)).to(torch.bfloat16)
    y = (O.bilinear_resize(base, (h2, w2)) + 0.4 * torch.randn(b, C, h2, w2, generator=g)).to(torch.bfloat16)
    pw = torch.randn(D, 2, generator=g) * 0.5; pb = torch.randn(D, generator=g) * 0.5
    want = O.gp_posterior(x.float(), y.float(), pw, pb, dtype=torch.float64)
    ref32 = O.gp_posterior(x.float(), y.float(), pw, pb)
    xc, yc, pwc, pbc = x.to(cuda), y.to(cuda), pw.to(cuda), pb.to(cuda)
    try:
        assert lib.roma_gp_set_path(3) == 0
        assert lib.roma_gp_set_lanes(0) == 0
        single = ops.gp_posterior(xc, yc, pwc, pbc)
        assert lib.roma_gp_set_lanes(1) == 0
        lanes = ops.gp_posterior(xc, yc, pwc, pbc)
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        reps = []
        with torch.cuda.stream(side):
            for _ in range(4):
                reps.append(ops.gp_posterior(xc, yc, pwc, pbc))
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
    finally:
        lib.roma_gp_set_lanes(1)
        lib.roma_gp_set_path(0)
    e_l, e_s, e_ref = err(lanes, want), err(single, want), err(ref32, want)
    assert e_l < max(10 * e_ref, 2e-4), (e_l, e_s, e_ref)
    assert err(lanes, single) < 2e-4
    for r in reps:
        assert torch.equal(r, lanes)


@pytest.mark.parametrize("dtype,split,tol", [(torch.float32, 3, 2e-6), (torch.float32, 2, 2e-5), (torch.bfloat16, 1, 2e-6), (torch.float16, 1, 2e-6)])
@pytest.mark.parametrize("batch,M,N,K", [(2, 300, 200, 136), (1, 128, 128, 64), (3, 129, 257, 520)])
def test_tc_gemm_nt(ops, cuda, dtype, split, tol, batch, M, N, K):
    """The tcgen05 building block on ragged shapes (row/column/K tails) against fp64.  16-bit operands run the
    TMA-fed warp-specialised main loop (default) and the register-staged one: identical results."""
    from roma_b200 import _lib
    g = torch.Generator().manual_seed(5)
    a = torch.randn(batch, M, K, generator=g).to(dtype); b = torch.randn(batch, N, K, generator=g).to(dtype)
    want = a.double() @ b.double().mT
    got = ops.tc_gemm_nt(a.to(cuda), b.to(cuda), split)
    assert (got.double().cpu() - want).abs().max().item() < tol * want.abs().max().item()
    if dtype != torch.float32:
        try:
            _lib.load().roma_tc_set_tma(0)
            staged = ops.tc_gemm_nt(a.to(cuda), b.to(cuda), split)
        finally:
            _lib.load().roma_tc_set_tma(1)
        assert torch.equal(staged, got)


# --------------------------------------------------------------------------------------------- K3b
@pytest.mark.parametrize("N,down", [(5000, 1), (40000, 1), (24001, 8), (1023, 1), (37, 3)])
@pytest.mark.parametrize("mode", [0, 2, 4])
def test_kde_lut_equals_direct(ops, cuda, N, down, mode):
    """The table-driven half paths against evaluating the reference's half op chain per pair (mode 1).
    mode 2 (SIMT FFMA2 distances) contributes the identical half value for every pair; modes 0 (default, exact: see
    tests/test_kde_exact_gpu.py) and 4 (fast, every pair once) take the distances from one tensor-core mma.sync per
    16 x 8 pairs, like the reference's HGEMM, which sums the same six exact products in a different fp32 order than the
    sequential chain, so a pair can fall on the neighbouring half value of d2.  Either way only the final half rounding
    of a few rows may move, by one ulp."""
    from roma_b200 import _lib
    lib = _lib.load()
    g = torch.Generator().manual_seed(9)
    centres = torch.rand(12, 4, generator=g) * 1.6 - 0.8
    x = (centres[torch.randint(0, 12, (N,), generator=g)] + 0.08 * torch.randn(N, 4, generator=g)).clamp(-1, 1).to(cuda)
    try:
        lib.roma_kde_set_direct(1)
        d_direct, p_direct = ops.kde(x, half=True, down=down, return_p=True)
        lib.roma_kde_set_direct(mode)
        d_lut, p_lut = ops.kde(x, half=True, down=down, return_p=True)
    finally:
        lib.roma_kde_set_direct(0)
    if N >= 5000:
        assert d_direct.float().max().item() > 50          # a non-trivial density field
    diff = (d_direct.float() - d_lut.float()).abs()
    ulp = d_direct.float().abs().clamp(min=1.0) * 2.0 ** -10
    assert (diff <= ulp).all() and (diff > 0).float().mean().item() < (2e-3 if mode == 2 else 1e-2)


@pytest.mark.parametrize("N", [40000, 5003, 1023, 130, 37])
def test_kde_symmetric_evaluation(ops, cuda, N):
    """Mode 4 evaluates every pair once (row sum of i + column sum of j); mode 3 evaluates (i, j) and (j, i): the same
    terms in another fp32 order, so the half densities agree up to the final rounding of a few rows; the symmetric
    path is deterministic run to run (no atomics) and permutation-covariant."""
    from roma_b200 import _lib
    lib = _lib.load()
    g = torch.Generator().manual_seed(N)
    centres = torch.rand(9, 4, generator=g) * 1.6 - 0.8
    x = (centres[torch.randint(0, 9, (N,), generator=g)] + 0.06 * torch.randn(N, 4, generator=g)).clamp(-1, 1).to(cuda)
    try:
        lib.roma_kde_set_direct(3)
        d_full = ops.kde(x, half=True).float()
        lib.roma_kde_set_direct(4)
        d_sym = ops.kde(x, half=True).float()
        d_sym2 = ops.kde(x, half=True).float()
        perm = torch.randperm(N, generator=g).to(cuda)
        d_perm = ops.kde(x[perm], half=True).float()
    finally:
        lib.roma_kde_set_direct(0)
    assert torch.equal(d_sym, d_sym2)
    diff = (d_full - d_sym).abs()
    assert (diff <= d_full.clamp(min=1.0) * 2.0 ** -10).all() and (diff > 0).float().mean().item() < 1e-2
    if N >= 1000:
        assert d_full.max().item() > 20
    dp = (d_perm - d_sym[perm]).abs()
    assert (dp <= d_sym[perm].clamp(min=1.0) * 2.0 ** -10).all()


def test_kde_tensor_core_distances_fp32_sums(ops, cuda):
    """Before the final half rounding: the fp32 row sums of the tensor-core path and of the direct chain agree to
    1e-5 relative (a pair moving to the neighbouring table entry changes its term by ~1e-3 of itself)."""
    from roma_b200 import _lib
    lib = _lib.load()
    g = torch.Generator().manual_seed(19)
    N = 20000
    x = (torch.rand(N, 4, generator=g) * 2 - 1)
    x[: N // 2] = x[N // 2:] * 0.1 + 0.3 * torch.randn(1, 4, generator=g)
    x = x.clamp(-1, 1).to(cuda)
    try:
        lib.roma_kde_set_direct(1)
        d1 = ops.kde(x, half=True).float()
        lib.roma_kde_set_direct(0)
        d0 = ops.kde(x, half=True).float()
    finally:
        lib.roma_kde_set_direct(0)
    rel = ((d0 - d1).abs() / d1.clamp(min=1.0)).max().item()
    assert rel <= 2.0 ** -10 and d1.max().item() > 1000


def test_kde_golden_and_oracle(ops, cuda, golden):
    g = golden("kde")
    d32 = ops.kde(g["x"].to(cuda), half=False)
    # fp32: the kernel sums (a - b)^2 directly; the reference's cdist forms |a|^2 + |b|^2 - 2ab in fp32, whose cancellation
    # error (~4e-7 absolute in d2) is ~2e-5 relative in exp(-d2 / 0.02).  That, not the kernel, sets the tolerance.
    assert ((d32.cpu() - g["dens_f32"]).abs() / g["dens_f32"]).max().item() < 1e-4
    dd = ops.kde(g["x"].to(cuda), half=False, down=8)
    assert ((dd.cpu() - g["dens_down"]).abs() / g["dens_down"].clamp(min=1e-3)).max().item() < 1e-4
    d16 = ops.kde(g["x"].to(cuda), half=True)
    assert d16.dtype == torch.float16
    rel = ((d16.float().cpu() - g["dens_half"]).abs() / g["dens_half"].clamp(min=1)).max().item()
    assert rel < 4e-3          # within a few fp16 ulps of the reference's CPU fp16 chain


def test_kde_weights(ops, cuda):
    g = torch.Generator().manual_seed(2)
    x = torch.cat((torch.rand(3000, 4, generator=g) * 2 - 1, torch.randn(1000, 4, generator=g) * 0.02)).to(cuda)
    dens, p = ops.kde(x, half=False, return_p=True)
    want = 1 / (dens + 1)
    want[dens < 10] = 1e-7
    assert torch.equal(p, want)


def test_sample_matches_oracle_given_same_density(ops, cuda):
    """With the kde swapped for the kernel inside the oracle's sample(), both consume the SAME CUDA generator
    stream: the sampled flat indices must be bit-identical."""
    g = torch.Generator().manual_seed(8)
    H, W = 64, 96
    warp = torch.rand(H, W, 4, generator=g).to(cuda) * 2 - 1
    cert = (torch.rand(H, W, generator=g) * (torch.rand(H, W, generator=g) < 0.6)).to(cuda)
    gen = torch.Generator(device=cuda)
    gen.manual_seed(1234)
    m1, c1, i1 = ops.sample(warp, cert, num=500, generator=gen, return_indices=True)
    gen.manual_seed(1234)
    m0, c0, i0 = O.sample(warp, cert, num=500, generator=gen, kde_fn=lambda x, std: ops.kde(x, std=std), return_indices=True)
    assert i1.dtype == torch.int64 and torch.equal(i1, i0) and torch.equal(m1, m0) and torch.equal(c1, c0)


# --------------------------------------------------------------------------------------------- K3c
def test_match_keypoints_golden(ops, cuda, golden):
    g = golden("match_keypoints")
    ia, ib = ops.match_keypoints(g["xA"].to(cuda), g["xB"].to(cuda), g["warp"].to(cuda), g["cert"].to(cuda))
    assert ia.dtype == torch.int64
    assert torch.equal(ia.cpu(), g["inds_A"]) and torch.equal(ib.cpu(), g["inds_B"])


def test_match_keypoints_vs_oracle_large(ops, cuda):
    g = torch.Generator().manual_seed(13)
    H, W = 60, 160
    warp = torch.cat((O.pixel_grid(1, H, W), make_flow(g, 1, H, W, "coherent")), dim=1)[0].permute(1, 2, 0).contiguous()
    cert = torch.rand(H, W, generator=g)
    xA = torch.rand(1500, 2, generator=g) * 2 - 1
    xB = torch.rand(1700, 2, generator=g) * 2 - 1
    ia0, ib0 = O.match_keypoints(xA, xB, warp, cert)
    ia1, ib1 = ops.match_keypoints(xA.to(cuda), xB.to(cuda), warp.to(cuda), cert.to(cuda))
    assert torch.equal(ia1.cpu(), ia0) and torch.equal(ib1.cpu(), ib0)


# ------------------------------------------------------------------------------------- f-1, Tiny
def test_cls_to_flow_golden_and_strided(ops, cuda, golden):
    g = golden("cls_to_flow")
    assert err(ops.cls_to_flow_refine(g["cls"].to(cuda)), g["flow"]) < 1e-5
    # token-major logits with a trailing certainty column, as TransformerDecoder produces them
    gg = torch.Generator().manual_seed(4)
    out = torch.randn(2, 6 * 7, 4097, generator=gg) * 2
    cls = out.permute(0, 2, 1).reshape(2, 4097, 6, 7)[:, :-1]
    want = O.cls_to_flow_refine(cls)
    got = ops.cls_to_flow_refine(out.to(cuda).permute(0, 2, 1).reshape(2, 4097, 6, 7)[:, :-1])
    assert err(got, want) < 1e-5
    got16 = ops.cls_to_flow_refine(out.half().to(cuda).permute(0, 2, 1).reshape(2, 4097, 6, 7)[:, :-1])
    assert err(got16, O.cls_to_flow_refine(out.half().float().permute(0, 2, 1).reshape(2, 4097, 6, 7)[:, :-1])) < 1e-5


@pytest.mark.parametrize("tc", [True, False])
def test_tiny_golden_and_oracle(ops, cuda, golden, tc):
    """Fused corr_volume + pos_embed: the tcgen05 GEMM with the fused arg-max / online-softmax epilogue (tc=True,
    fp32 features as three bf16 planes) and the fp32 SIMT kernel, against the reference's own output and the oracle."""
    g = golden("tiny")
    assert err(ops.tiny_corr_posembed(g["feat0"].to(cuda), g["feat1"].to(cuda), tensor_cores=tc), g["pos_embed"]) < 1e-5
    gg = torch.Generator().manual_seed(6)
    for (h0, w0, h1, w1) in [(12, 16, 12, 16), (9, 15, 20, 28)]:          # ragged row / column tiles
        f0 = torch.randn(2, 64, h0, w0, generator=gg); f1 = torch.randn(2, 64, h1, w1, generator=gg)
        want = torch.cat([O.tiny_pos_embed(O.tiny_corr_volume(f0[i:i + 1], f1[i:i + 1])) for i in range(2)])
        assert err(ops.tiny_corr_posembed(f0.to(cuda), f1.to(cuda), tensor_cores=tc), want) < 1e-5


def test_tiny_corr_tensor_cores_config5_shape(ops, cuda):
    """BASELINE config 5 shape (60 x 80 = 4800 tokens per image, C = 64): the tensor-core path against the SIMT kernel.
    Outputs agree to 1e-5 wherever the two pick the same arg-max; a different pick needs two scores within fp32
    rounding of each other, so at most a handful of the 14400 queries may differ."""
    g = torch.Generator().manual_seed(12)
    f0 = torch.randn(3, 64, 60, 80, generator=g).to(cuda)
    f1 = (0.5 * f0.roll(3, 3) + torch.randn(3, 64, 60, 80, generator=g).to(cuda))
    a = ops.tiny_corr_posembed(f0, f1, tensor_cores=True)
    b = ops.tiny_corr_posembed(f0, f1, tensor_cores=False)
    assert a.shape == b.shape == (3, 2, 60, 80) and torch.isfinite(a).all()
    bad = ((a - b).abs().amax(dim=1) > 1e-5).float().mean().item()
    assert bad < 1e-3, bad


# --------------------------------------------------------------------------------------------- (f-2)
@pytest.mark.parametrize("dtype,tol", [(torch.float32, 1e-5), (torch.bfloat16, 2e-2), (torch.float16, 3e-3)])
@pytest.mark.parametrize("B,C,h,w", [(2, 24, 17, 23), (1, 144, 9, 8), (2, 576, 5, 13), (1, 2, 40, 3), (1, 1144, 33, 21),
                                     (1, 40, 16, 64), (3, 72, 31, 4), (1, 10, 37, 41),
                                     # strips of 36 rows inside the image with w % 4 == 0: the branch-free store path, at the
                                     # four compile-time refiner widths and a runtime one, each followed by a partial strip
                                     (1, 144, 80, 16), (1, 576, 40, 8), (1, 1144, 37, 12), (1, 1384, 72, 4), (2, 40, 75, 20)])
@pytest.mark.parametrize("path", [0, 1])
def test_dwconv5x5_bias_relu(ops, cuda, dtype, tol, B, C, h, w, path):
    """Depthwise 5x5 + bias + ReLU vs F.conv2d(groups=C) + relu in fp32 on the same (rounded) inputs; path 0 takes
    the TMA-tiled kernel where the channel stride allows it, path 1 forces the cp.async kernel."""
    from roma_b200 import _lib
    assert _lib.load().roma_dwconv_set_path(path) == 0
    try:
        _dwconv_case(ops, cuda, dtype, tol, B, C, h, w)
    finally:
        _lib.load().roma_dwconv_set_path(0)


def _dwconv_case(ops, cuda, dtype, tol, B, C, h, w):
    g = torch.Generator().manual_seed(17)
    x = torch.randn(B, C, h, w, generator=g).to(dtype)
    wt = torch.randn(C, 1, 5, 5, generator=g) * 0.2
    bias = torch.randn(C, generator=g)
    want = torch.relu(torch.nn.functional.conv2d(x.float(), wt, bias, padding=2, groups=C))
    taps = wt.view(C, 25).t().contiguous()
    got = ops.dwconv5x5_bias_act(x.to(cuda), taps.to(cuda), bias.to(cuda), relu=True)
    assert got.dtype == dtype and got.shape == want.shape
    assert err(got, want) < tol * max(1.0, want.abs().max().item())
    lin = ops.dwconv5x5_bias_act(x.to(cuda), taps.to(cuda), bias.to(cuda), relu=False)
    assert err(lin, torch.nn.functional.conv2d(x.float(), wt, bias, padding=2, groups=C)) < tol * max(1.0, want.abs().max().item())


@pytest.mark.parametrize("dtype,tol", [(torch.float32, 2e-5), (torch.bfloat16, 2e-5), (torch.float16, 2e-5)])
@pytest.mark.parametrize("B,C,h,w", [(2, 24, 17, 23), (1, 144, 9, 8), (2, 576, 5, 13), (1, 1144, 11, 7), (1, 8, 3, 5), (1, 64, 1, 1), (1, 1384, 6, 5),
                                     (1, 2560, 3, 3)])
def test_refiner_out_conv(ops, cuda, dtype, tol, B, C, h, w):
    """out_conv (C -> 3) in fp32 from the stored activation vs F.conv2d on the up-converted tensor (matcher.py:141)."""
    g = torch.Generator().manual_seed(23)
    x = torch.randn(B, C, h, w, generator=g).to(dtype)
    wt = torch.randn(3, C, generator=g) / C ** 0.5
    bias = torch.randn(3, generator=g)
    want = torch.nn.functional.conv2d(x.double(), wt.double().view(3, C, 1, 1), bias.double()).float()
    got = ops.refiner_out_conv(x.to(cuda).contiguous(memory_format=torch.channels_last), wt.to(cuda), bias.to(cuda))
    assert got.shape == want.shape and got.dtype == torch.float32
    assert err(got, want) < tol * max(1.0, want.abs().max().item())
    one = ops.refiner_out_conv(x.to(cuda), wt[:1].to(cuda), bias[:1].to(cuda))
    assert err(one, want[:, :1]) < tol * max(1.0, want.abs().max().item())


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.float16])
@pytest.mark.parametrize("B,C,h,w", [(2, 64, 18, 22), (1, 128, 7, 9), (3, 8, 2, 2), (1, 512, 5, 4)])
def test_maxpool2x2_bit_exact(ops, cuda, dtype, B, C, h, w):
    """VGG19 stage pooling (encoders.py:57-66): bit-identical to F.max_pool2d(2, 2), odd row / column dropped."""
    x = torch.randn(B, C, h, w, generator=torch.Generator().manual_seed(5)).to(dtype)
    want = torch.nn.functional.max_pool2d(x.float(), 2, 2).to(dtype)
    got = ops.maxpool2x2(x.to(cuda).contiguous(memory_format=torch.channels_last))
    assert got.shape == want.shape and got.dtype == dtype
    assert torch.equal(got.cpu(), want)


@pytest.mark.parametrize("dtype,tol", [(torch.float32, 2e-5), (torch.bfloat16, 1e-2), (torch.float16, 2e-3)])
@pytest.mark.parametrize("B,C,OUT,h,w", [(2, 64, 9, 17, 23), (1, 8, 1, 3, 5), (1, 128, 16, 4, 4), (3, 64, 9, 1, 1),
                                         (2, 24, 24, 9, 11), (1, 32, 32, 5, 5), (1, 16, 20, 3, 3), (1, 48, 7, 6, 7),
                                         (1, 8, 3, 37, 1)])
def test_pointwise_narrow(ops, cuda, dtype, tol, B, C, OUT, h, w):
    """proj["1"] (Conv 64 -> 9 + folded BN) as a streaming kernel vs F.conv2d in fp64 on the stored inputs."""
    g = torch.Generator().manual_seed(29)
    x = torch.randn(B, C, h, w, generator=g).to(dtype)
    wt = torch.randn(OUT, C, generator=g) / C ** 0.5
    bias = torch.randn(OUT, generator=g)
    want = torch.nn.functional.conv2d(x.double(), wt.double().view(OUT, C, 1, 1), bias.double()).float()
    from roma_b200 import _lib
    xc = x.to(cuda).contiguous(memory_format=torch.channels_last)
    got = ops.pointwise_narrow(xc, wt.to(cuda), bias.to(cuda))          # 16-bit: tensor-core path (weights in dtype)
    assert got.shape == want.shape and got.dtype == dtype
    assert err(got, want) < tol * max(1.0, want.abs().max().item())
    try:
        _lib.load().roma_pointwise_narrow_set_simt(1)
        simt = ops.pointwise_narrow(xc, wt.to(cuda), bias.to(cuda))     # fp32-weight FFMA2 kernel
    finally:
        _lib.load().roma_pointwise_narrow_set_simt(0)
    assert err(simt, want) < tol * max(1.0, want.abs().max().item())


@pytest.mark.parametrize("dtype,tol", [(torch.float32, 2e-6), (torch.bfloat16, 8e-3), (torch.float16, 1e-3)])
@pytest.mark.parametrize("shape", [(3, 17, 1024), (5, 64), (2, 7, 384), (1, 2048), (4, 8)])
def test_layer_norm(ops, cuda, dtype, tol, shape):
    """LayerNorm vs F.layer_norm evaluated in fp64 on the stored inputs (block.py:88-101)."""
    g = torch.Generator().manual_seed(31)
    x = (torch.randn(*shape, generator=g) * 3 + 0.5).to(dtype)
    D = shape[-1]
    wt = (1 + 0.1 * torch.randn(D, generator=g)).to(dtype)
    bias = (0.1 * torch.randn(D, generator=g)).to(dtype)
    if dtype == torch.float32 and D > 1024:
        pytest.skip("fp32 rows are limited to 1024 channels")
    want = torch.nn.functional.layer_norm(x.double(), (D,), wt.double(), bias.double(), 1e-6).float()
    got = ops.layer_norm(x.to(cuda), wt.to(cuda), bias.to(cuda), 1e-6)
    assert got.shape == x.shape and got.dtype == dtype
    assert err(got, want) < tol * max(1.0, want.abs().max().item())


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("B,h,w", [(2, 16, 40), (1, 37, 53), (3, 70, 81), (1, 5, 3), (2, 48, 200)])
def test_dwconv_pointwise_fused_block(ops, cuda, dtype, B, h, w):
    """The fused refiner block (C = 24): bit-identical to the depthwise kernel followed by the narrow 1x1 kernel (same
    rounding points, same MMA sequence), and within 16-bit rounding of an fp64 evaluation of the two convolutions."""
    g = torch.Generator().manual_seed(41)
    C = 24
    x = torch.randn(B, C, h, w, generator=g).to(dtype).to(cuda).contiguous(memory_format=torch.channels_last)
    w_dw = (torch.randn(25, C, generator=g) * 0.2).to(cuda)
    b_dw = (torch.randn(C, generator=g) * 0.1).to(cuda)
    w_pw = (torch.randn(C, C, generator=g) / C ** 0.5).to(cuda)
    b_pw = (torch.randn(C, generator=g) * 0.1).to(cuda)
    two = ops.pointwise_narrow(ops.dwconv5x5_bias_act(x, w_dw, b_dw, relu=True), w_pw, b_pw)
    one = ops.dwconv5x5_pointwise(x, w_dw, b_dw, w_pw, b_pw, relu=True)
    assert one.shape == two.shape == (B, C, h, w) and one.dtype == dtype
    assert torch.equal(one, two)
    mid = torch.nn.functional.conv2d(x.double(), w_dw.t().reshape(C, 1, 5, 5).double(), b_dw.double(), padding=2, groups=C).relu()
    mid = mid.to(dtype).double()                                   # the 16-bit activation between the two convolutions
    want = torch.nn.functional.conv2d(mid, w_pw.double().view(C, C, 1, 1), b_pw.double())
    tol = (2.0 ** -9 if dtype == torch.float16 else 2.0 ** -6) * max(1.0, want.abs().max().item())
    assert (one.double() - want).abs().max().item() < tol


def test_dwconv_pointwise_rejects_other_widths(ops, cuda):
    x = torch.zeros(1, 32, 8, 8, device=cuda, dtype=torch.bfloat16).contiguous(memory_format=torch.channels_last)
    with pytest.raises((RuntimeError, ValueError)):
        ops.dwconv5x5_pointwise(x, torch.zeros(25, 32, device=cuda), torch.zeros(32, device=cuda), torch.zeros(32, 32, device=cuda),
                                torch.zeros(32, device=cuda))


# --------------------------------------------------------------------------------------------- (f-3)
@pytest.mark.parametrize("H,W,h,w", [(864, 864, 560, 560), (480, 640, 560, 560), (600, 1200, 224, 308), (64, 48, 64, 48),
                                     (100, 75, 333, 290), (1120, 896, 280, 224), (63, 47, 63, 47), (35, 29, 17, 23)])
def test_prepare_images_vs_aten(ops, cuda, H, W, h, w):
    """uint8 batch -> normalised (+ antialiased bicubic resize): against (u8/255 - mean)/std followed by
    F.interpolate(antialias=True).  The reference pipeline runs that on the GPU, and the kernel reproduces ATen's CUDA
    arithmetic (weights, normalisation, summation order): compared with ATen on the same device; ATen's CPU kernel
    (different vectorisation, itself up to 1e-4 away from ATen CUDA) is the looser cross-check.  Same size: bit-identical."""
    import torch.nn.functional as F
    g = torch.Generator().manual_seed(H * 7 + w)
    img = (torch.rand(2, 3, H, W, generator=g) * 255).round().to(torch.uint8)
    mean = torch.tensor([0.485, 0.456, 0.406]).view(1, 3, 1, 1)
    std = torch.tensor([0.229, 0.224, 0.225]).view(1, 3, 1, 1)
    norm = (img.float() / 255.0 - mean) / std
    got = ops.prepare_images(img.to(cuda), (h, w))
    assert got.shape == (2, 3, h, w) and got.dtype == torch.float32
    if (h, w) == (H, W):
        assert torch.equal(got.cpu(), norm)
    else:
        norm_dev = (img.to(cuda).float() / 255.0 - mean.to(cuda)) / std.to(cuda)
        want_dev = F.interpolate(norm_dev, size=(h, w), mode="bicubic", antialias=True, align_corners=False)
        want_cpu = F.interpolate(norm, size=(h, w), mode="bicubic", antialias=True, align_corners=False)
        print(f"prepare_images {H}x{W}->{h}x{w}: vs ATen CUDA {err(got, want_dev):.3g}, vs ATen CPU {err(got, want_cpu):.3g}")
        assert err(got, want_dev) <= 2e-6
        assert err(got, want_cpu) < 5e-4


PIL_CASES = [((48, 64), (35, 40)), ((480, 640), (560, 560)), ((618, 640), (864, 864)), ((100, 75), (333, 290)),
             ((600, 1200), (224, 308)), ((37, 53), (37, 80)), ((64, 48), (20, 48)), ((30, 30), (30, 30)),
             ((1120, 896), (280, 224)), ((5, 7), (16, 3))]


@pytest.mark.parametrize("src,dst", PIL_CASES)
@pytest.mark.parametrize("kind", ["noise", "edges"])
def test_prepare_images_pil_bit_exact(ops, cuda, src, dst, kind):
    """PIL bicubic resize of 8-bit images on the device: equal, byte for byte, to Pillow itself and to the restatement
    in oracle/pil_resample_oracle.py; interleaved (as decoded) and planar inputs; the normalised fp32 output equals
    the reference's ToTensor + Normalize applied to Pillow's result, bit for bit."""
    import numpy as np
    from PIL import Image
    from oracle import pil_resample_oracle as P
    g = np.random.default_rng(src[0] * 31 + dst[1])
    if kind == "noise":
        imgs = [g.integers(0, 256, size=(*src, 3), dtype=np.uint8) for _ in range(2)]
    else:
        imgs = [np.kron(g.integers(0, 2, size=(src[0] // 4 + 1, src[1] // 4 + 1, 3)) * 255,
                        np.ones((4, 4, 1)))[: src[0], : src[1]].astype(np.uint8) for _ in range(2)]
    want = np.stack([np.asarray(Image.fromarray(im).resize((dst[1], dst[0]), resample=Image.BICUBIC)) for im in imgs])
    assert np.array_equal(want[0], P.resize_bicubic_u8(imgs[0], dst))
    hwc = torch.from_numpy(np.stack(imgs)).to(cuda)
    got = ops.prepare_images_pil(hwc, dst, channels_last=True, normalise=False)
    assert got.shape == (2, 3, *dst) and got.dtype == torch.uint8
    assert np.array_equal(got.permute(0, 2, 3, 1).cpu().numpy(), want)
    planar = hwc.permute(0, 3, 1, 2).contiguous()
    assert torch.equal(ops.prepare_images_pil(planar, dst, normalise=False), got)
    mean = torch.tensor([0.485, 0.456, 0.406]).view(1, 3, 1, 1)
    std = torch.tensor([0.229, 0.224, 0.225]).view(1, 3, 1, 1)
    ref = (torch.from_numpy(want.astype(np.float32).transpose(0, 3, 1, 2) / 255.0) - mean) / std      # utils.py:234-261
    assert torch.equal(ops.prepare_images_pil(hwc, dst, channels_last=True).cpu(), ref)


def test_prepare_images_pil_rejects(ops, cuda):
    with pytest.raises((RuntimeError, ValueError)):
        ops.prepare_images_pil(torch.zeros(1, 3, 900, 900, dtype=torch.uint8, device=cuda), (200, 200))
    with pytest.raises((RuntimeError, ValueError)):
        ops.prepare_images_pil(torch.zeros(1, 3, 90, 90, device=cuda), (20, 20))


def test_prepare_images_rejects_large_downscale(ops, cuda):
    with pytest.raises((RuntimeError, ValueError)):
        ops.prepare_images(torch.zeros(1, 3, 900, 900, dtype=torch.uint8, device=cuda), (200, 200))


# --------------------------------------------------------------------------------------------- encoder side
@pytest.mark.parametrize("dtype,tol", [(torch.float32, 2e-5), (torch.bfloat16, 3e-2), (torch.float16, 4e-3)])
def test_layernorm_shift(ops, cuda, dtype, tol):
    """LayerNorm(x + shift) with the shift added in fp32 inside the kernel vs F.layer_norm on the fp32 sum."""
    g = torch.Generator().manual_seed(17)
    x = (torch.randn(3, 37, 1024, generator=g) * 2 + 0.5).to(dtype)
    shift = torch.randn(1024, generator=g) * 0.7
    w = (1 + 0.1 * torch.randn(1024, generator=g)).to(dtype); b = (0.1 * torch.randn(1024, generator=g)).to(dtype)
    want = torch.nn.functional.layer_norm(x.float() + shift, (1024,), w.float(), b.float(), 1e-6)
    got = ops.layer_norm(x.to(cuda), w.to(cuda), b.to(cuda), 1e-6, shift=shift.to(cuda))
    assert got.dtype == dtype and err(got, want) < tol
    assert torch.equal(ops.layer_norm(x.to(cuda), w.to(cuda), b.to(cuda), 1e-6, shift=torch.zeros(1024, device=cuda)),
                       ops.layer_norm(x.to(cuda), w.to(cuda), b.to(cuda), 1e-6))


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.float16])
def test_gelu_erf(ops, cuda, dtype):
    """roma_gelu against the exact GELU x * Phi(x) in fp64 (erfc form): fp32 to 4e-7 absolute and 1e-5 relative
    including the negative tail; 16-bit outputs equal to the correctly rounded exact value; at most one unit in the
    last place from ATen's GELU; in place == out of place; special values."""
    g = torch.Generator().manual_seed(5)
    x = torch.cat([torch.randn(1 << 20, generator=g) * 2.5, torch.linspace(-9, 9, 1 << 16),
                   torch.tensor([0.0, -0.0, 1e-30, -1e-30, 30.0, -30.0, 1e4, -1e4])]).to(dtype)
    xd = x.double()
    exact = 0.5 * xd * torch.erfc(-xd / 2 ** 0.5)
    got = ops.gelu(x.to(cuda))
    assert got.dtype == dtype and got.shape == x.shape
    if dtype == torch.float32:
        d = (got.cpu().double() - exact).abs()
        assert d.max().item() < 4e-7
        m = (exact.abs() > 1e-30) & (xd.abs() < 6)
        assert (d[m] / exact[m].abs()).max().item() < 1e-5                       # relative accuracy in the negative tail
        m = (exact.abs() > 1e-30) & (xd.abs() < 9.01)
        assert (d[m] / exact[m].abs()).max().item() < 5e-5                       # exp argument ~ -60: its own rounding
    else:
        assert (got.cpu() != exact.to(dtype)).float().mean().item() < 1e-4          # correctly rounded almost everywhere
        aten = torch.nn.functional.gelu(x.to(cuda))
        ulp = torch.finfo(dtype).eps * aten.float().abs().clamp_min(torch.finfo(dtype).tiny)
        assert ((got.float() - aten.float()).abs() <= 1.01 * ulp + 5e-7).all()        # 5e-7: ATen's cancellation in 1 + erf
    assert torch.isfinite(got).all()
    y = x.to(cuda).clone()
    assert ops.gelu(y, inplace=True).data_ptr() == y.data_ptr() and torch.equal(y, got)
    with pytest.raises((RuntimeError, ValueError)):
        ops.gelu(torch.zeros(7, device=cuda, dtype=dtype))


@pytest.mark.parametrize("dtype,tol", [(torch.float32, 1e-4), (torch.bfloat16, 6e-2)])
def test_vit_unbiased_residual_stream(cuda, dtype, tol):
    """DinoViT with the residual branches as in-place accumulating GEMMs (biases carried as an fp32 shift into the
    LayerNorms) vs the plain block formulation, same weights with NON-zero biases and LayerScale."""
    from roma_b200.host.vit import DinoViT
    torch.manual_seed(3)
    vit = DinoViT(img_size=70, patch_size=14, embed_dim=128, depth=3, num_heads=4, init_values=0.7)
    with torch.no_grad():
        for m in vit.modules():
            if isinstance(m, torch.nn.Linear):
                m.bias.normal_(std=0.3)
    vit = vit.to(cuda).to(dtype).eval()
    x = torch.randn(2, 3, 70, 84, device=cuda).to(dtype)
    with torch.inference_mode():
        vit.unbiased_residual = True
        a = vit.patch_tokens(x)
        vit.unbiased_residual = False
        b = vit.patch_tokens(x)
    assert a.shape == b.shape == (2, 30, 128)
    assert err(a, b) < tol


def test_layernorm_cast_and_residual_add(ops, cuda):
    """fp32 rows -> 16-bit LayerNorm output = F.layer_norm in fp32 followed by the cast autocast inserts (same bits);
    fp32 + 16-bit residual add = ATen's mixed-dtype add (same bits)."""
    g = torch.Generator(device="cuda").manual_seed(21)
    x = torch.randn(3, 257, 1024, device=cuda, generator=g) * 3 + 0.5
    w = torch.randn(1024, device=cuda, generator=g)
    b = torch.randn(1024, device=cuda, generator=g)
    for dt in (torch.bfloat16, torch.float16):
        want = torch.nn.functional.layer_norm(x, (1024,), w, b, 1e-5)
        got = ops.layer_norm(x, w, b, 1e-5, out_dtype=dt)
        assert got.dtype == dt
        # the fp32 statistics differ in the last bits (two-pass vs Welford): the 16-bit roundings agree except at ties
        assert (got.float() - want.to(dt).float()).abs().max().item() <= 2.0 ** (-7 if dt == torch.bfloat16 else -10) * want.abs().max().item()
        assert (got != want.to(dt)).float().mean().item() < 2e-3
        y = torch.randn(3, 257, 1024, device=cuda, generator=g).to(dt)
        assert torch.equal(ops.residual_add(x, y), x + y)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
@pytest.mark.parametrize("rows,D", [(1601 * 2, 1024), (77, 384), (5, 8), (33, 768), (9, 256), (3, 2048), (11, 512), (7, 128)])
def test_add_layernorm_is_add_then_layernorm(ops, cuda, dtype, rows, D):
    """roma_add_layernorm: the sum is ATen's x + y bit for bit, the norm is roma_layernorm of that sum bit for bit; in place too."""
    if dtype == torch.float32 and D > 1024:
        pytest.skip("fp32 rows: D <= 1024")
    g = torch.Generator().manual_seed(rows + D)
    x = (torch.randn(rows, D, generator=g) * 3).to(dtype).to(cuda)
    y = torch.randn(rows, D, generator=g).to(dtype).to(cuda)
    w, b = (torch.rand(D, generator=g) + 0.5).to(dtype).to(cuda), torch.randn(D, generator=g).to(dtype).to(cuda)
    want_sum = x + y
    want_norm = ops.layer_norm(want_sum, w, b, 1e-6)
    s, n = ops.add_layer_norm(x, y, w, b, 1e-6)
    assert torch.equal(s, want_sum) and torch.equal(n, want_norm)
    xi = x.clone()
    s2, n2 = ops.add_layer_norm(xi, y, w, b, 1e-6, inplace=True)
    assert s2.data_ptr() == xi.data_ptr() and torch.equal(xi, want_sum) and torch.equal(n2, want_norm)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_vit_fused_residual_norm_is_bit_identical(cuda, dtype):
    """DinoViT with the residual adds fused into the following norms returns the same bits as the plain block loop."""
    from roma_b200.host.vit import DinoViT
    torch.manual_seed(3)
    vit = DinoViT(img_size=70, patch_size=14, embed_dim=256, depth=3, num_heads=4).to(cuda).to(dtype).eval()
    with torch.no_grad():
        for p in vit.parameters():
            if p.dim() == 1:
                p.add_(torch.randn_like(p) * 0.1)
    x = torch.randn(2, 3, 70, 84, device=cuda, dtype=dtype)
    with torch.inference_mode():
        vit.fused_residual_norm = True
        a = vit.patch_tokens(x)
        vit.fused_residual_norm = False
        b = vit.patch_tokens(x)
    assert a.shape == b.shape == (2, 30, 256) and torch.equal(a, b)


@pytest.mark.parametrize("dtype,tol", [(torch.float32, 2e-5), (torch.bfloat16, 3e-2), (torch.float16, 4e-3)])
@pytest.mark.parametrize("D", [8, 128, 256, 384, 512, 768, 1024, 2048])
def test_layernorm_all_row_widths(ops, cuda, dtype, tol, D):
    """roma_layernorm vs F.layer_norm in fp32 over rows that fill the lanes' vectors exactly (predicate-free variant)
    and rows that do not; the row after the last one is poisoned to catch reads past a row."""
    if dtype == torch.float32 and D > 1024:
        pytest.skip("fp32 rows: D <= 1024")
    g = torch.Generator().manual_seed(D)
    x = (torch.randn(19, D, generator=g) * 2 + 0.3).to(dtype)
    w = (1 + 0.1 * torch.randn(D, generator=g)).to(dtype); b = (0.1 * torch.randn(D, generator=g)).to(dtype)
    want = torch.nn.functional.layer_norm(x.float(), (D,), w.float(), b.float(), 1e-6)
    buf = torch.full((20, D), float("nan"), dtype=dtype, device=cuda)
    buf[:19] = x.to(cuda)
    got = ops.layer_norm(buf[:19], w.to(cuda), b.to(cuda), 1e-6)
    assert got.dtype == dtype and torch.isfinite(got).all() and err(got, want) < tol
