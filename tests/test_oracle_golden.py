"""The oracle restatement vs the golden vectors produced by the REFERENCE (tests/golden/make_golden.py)."""
import pytest
import torch

from oracle import roma_oracle as O


def close(a, b, tol):
    assert a.shape == b.shape, (a.shape, b.shape)
    err = (a.float() - b.float()).abs().max().item()
    assert err <= tol, f"max abs err {err} > {tol}"


@pytest.mark.parametrize("tag", ["r2", "r3", "r7"])
def test_local_correlation_and_grid_sample(golden, tag):
    g = golden(f"local_corr_{tag}")
    close(O.local_correlation(g["f0"], g["f1"], g["r"], g["flow"]), g["corr"], 2e-5)
    close(O.grid_sample_zeros(g["f1"], g["flow"].permute(0, 2, 3, 1)), g["xhat"], 1e-5)


def test_local_correlation_aligned(golden):
    g = golden("local_corr_noflow")
    close(O.local_correlation(g["f0"], g["f1"], g["r"], None), g["corr"], 2e-5)


@pytest.mark.parametrize("tag", ["corr", "nocorr"])
def test_refiner_inputs(golden, tag):
    g = golden(f"refiner_inputs_{tag}")
    d = O.refiner_inputs(g["x"], g["y"], g["flow"], g["disp_w"], g["disp_b"], 40 / 32 * g["scale_factor"], g["r"] or None)
    close(d, g["d"], 3e-5)


def test_update_and_resample(golden):
    g = golden("update_resample")
    fu, cu = O.flow_update(g["flow"], g["cert"], g["delta"], g["ins"], g["Hf"], g["Wf"])
    assert torch.equal(fu, g["flow_u"]) and torch.equal(cu, g["cert_u"])
    for k in ("up", "odd", "down"):
        size = tuple(int(v) for v in g[f"size_{k}"])
        close(O.bilinear_resize(fu, size), g[f"flow_{k}"], 1e-5)
        close(O.bilinear_resize(cu, size), g[f"cert_{k}"], 1e-5)


@pytest.mark.parametrize("tag", ["sq", "rect"])
def test_gp(golden, tag):
    g = golden(f"gp_{tag}")
    xf = g["x"].flatten(2).transpose(1, 2)
    yf = g["y"].flatten(2).transpose(1, 2)
    close(O.cos_kernel(xf, yf), g["Kxy"], 1e-6)
    close(O.gp_posterior(g["x"], g["y"], g["pos_w"], g["pos_b"]), g["mu"], 2e-5)
    close(O.gp_posterior(g["x"], g["y"], g["pos_w"], g["pos_b"], dtype=torch.float64), g["mu"], 1e-4)


@pytest.mark.parametrize("tag", ["sym", "asym"])
def test_match_epilogue(golden, tag):
    g = golden(f"epilogue_{tag}")
    warp, cert, _ = O.match_epilogue(g["flow"], g["cert"], g["cert16"] if g["attenuate"] else None, bool(g["symmetric"]))
    close(warp[0], g["warp"], 1e-6)
    close(cert[0], g["certainty"], 1e-6)


def test_kde(golden):
    g = golden("kde")
    assert torch.equal(O.kde(g["x"], half=True).float(), g["dens_half"])       # fp16 chain is bit-exact on CPU
    close(O.kde(g["x"], half=False), g["dens_f32"], 2e-4)
    close(O.kde(g["x"], half=False, down=8), g["dens_down"], 1e-4)


def test_sample_same_rng_stream(golden):
    g = golden("sample")
    torch.manual_seed(g["seed"])
    m, c, idx = O.sample(g["warp"], g["cert"], num=g["num"], return_indices=True)
    assert torch.equal(m, g["matches"]) and torch.equal(c, g["certainty"])
    assert torch.equal(g["warp"].reshape(-1, 4)[idx], g["matches"])


def test_match_keypoints(golden):
    g = golden("match_keypoints")
    ia, ib = O.match_keypoints(g["xA"], g["xB"], g["warp"], g["cert"])
    assert torch.equal(ia, g["inds_A"]) and torch.equal(ib, g["inds_B"])


def test_cls_to_flow_refine(golden):
    g = golden("cls_to_flow")
    close(O.cls_to_flow_refine(g["cls"]), g["flow"], 1e-6)


def test_tiny(golden):
    g = golden("tiny")
    cv = O.tiny_corr_volume(g["feat0"], g["feat1"])
    close(cv, g["corr_volume"], 1e-5)
    close(O.tiny_pos_embed(g["corr_volume"]), g["pos_embed"], 1e-5)


def test_aten_mode_equals_explicit_restatement():
    """The ATen-call mode used for the timed CPU baseline computes the same thing as the explicit restatements."""
    import torch
    from oracle import roma_oracle as O
    g = torch.Generator().manual_seed(3)
    f0, f1 = torch.randn(2, 16, 9, 11, generator=g), torch.randn(2, 16, 9, 11, generator=g)
    flow = torch.rand(2, 2, 9, 11, generator=g) * 2.4 - 1.2
    pts = torch.rand(700, 4, generator=g) * 2 - 1
    a = (O.local_correlation(f0, f1, 2, flow), O.grid_sample_zeros(f1, flow.permute(0, 2, 3, 1)), O.kde(pts, half=False))
    O.set_aten(True)
    try:
        b = (O.local_correlation(f0, f1, 2, flow), O.grid_sample_zeros(f1, flow.permute(0, 2, 3, 1)), O.kde(pts, half=False))
    finally:
        O.set_aten(False)
    for u, v in zip(a, b):
        assert (u - v).abs().max().item() < 1e-4 * max(1.0, u.abs().max().item())
