"""Tiny-RoMa (BASELINE config 5): oracle vs the reference's own `TinyRoMa.match` output (CPU), and the CUDA
product model vs both (GPU).  Fixture: tests/golden/tiny_e2e.npz (tests/golden/make_golden_tiny.py)."""
import numpy as np
import pytest
import torch

import roma_b200
from roma_b200.host.tiny import TinyRoMa, XFeatBackbone, random_init_tiny
from oracle import roma_model_oracle as MO

WARP_TOL, LOGIT_TOL = 1e-3, 1e-2


def _build(device):
    weights, trunk_sd = random_init_tiny(seed=0)
    trunk = XFeatBackbone()
    trunk.load_state_dict(trunk_sd)
    model = TinyRoMa(trunk)
    model.load_state_dict(weights)
    digest = roma_b200.state_digest({**weights, **{"xfeat." + k: v for k, v in trunk_sd.items()}})
    return model.to(device).train(False), digest


def _images(g):
    t = lambda im: im.permute(2, 0, 1)[None].float() / 255
    return t(g["im_A"]), t(g["im_B"])


def test_state_dict_keys_are_the_references():
    """coarse_matcher.* / fine_matcher.* only: the frozen trunk is hidden from the state dict (tiny.py:107-109)."""
    sd, _ = random_init_tiny(seed=0)
    assert all(k.split(".")[0] in ("coarse_matcher", "fine_matcher") for k in sd)
    assert "coarse_matcher.4.weight" in sd and sd["coarse_matcher.0.layer.0.weight"].shape == (256, 130, 3, 3)
    assert sd["fine_matcher.0.layer.0.weight"].shape == (64, 50, 3, 3)


def test_oracle_tiny_match_vs_reference_golden(golden):
    g = golden("tiny_e2e")
    model, digest = _build("cpu")
    assert digest == bytes(g["digest"].numpy().astype(np.uint8)).decode()
    A, B = _images(g)
    with torch.inference_mode():
        warp, cert = MO.tiny_match(model, A, B)
        corr = MO.tiny_forward(model, A, B)
    assert (warp[:, ::2, ::2] - g["warp"]).abs().max().item() < 1e-5
    assert (cert[:, ::2, ::2] - g["certainty"]).abs().max().item() < 1e-5
    for s in (8, 4):
        assert (corr[s]["flow"] - g[f"flow{s}"]).abs().max().item() < 1e-5
        assert (corr[s]["certainty"] - g[f"cert{s}"]).abs().max().item() < 1e-4


def test_product_model_refuses_cpu(golden):
    g = golden("tiny_e2e")
    model, _ = _build("cpu")
    A, B = _images(g)
    with pytest.raises((RuntimeError, ValueError)):
        model.match(A, B)


@pytest.mark.gpu
def test_tiny_match_gpu_vs_golden_and_oracle(golden, cuda):
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    g = golden("tiny_e2e")
    model, _ = _build(cuda)
    A, B = _images(g)
    warp, cert = model.match(A.to(cuda), B.to(cuda), batched=True)
    assert warp.shape == (1, 200, 232, 4) and cert.shape == (1, 200, 232)
    assert (warp[:, ::2, ::2].cpu() - g["warp"]).abs().max().item() < WARP_TOL
    assert (cert[:, ::2, ::2].cpu() - g["certainty"]).abs().max().item() < LOGIT_TOL
    with torch.inference_mode():
        corr = model.forward({"im_A": A.to(cuda), "im_B": B.to(cuda)})
    for s in (8, 4):
        assert (corr[s]["flow"].cpu() - g[f"flow{s}"]).abs().max().item() < WARP_TOL, f"stride {s} flow"
        assert (corr[s]["certainty"].cpu() - g[f"cert{s}"]).abs().max().item() < LOGIT_TOL, f"stride {s} certainty"
    # batch of 2 (B<->A swapped in the second element) equals two single calls
    w2, c2 = model.match(torch.cat((A, B)).to(cuda), torch.cat((B, A)).to(cuda), batched=True)
    assert (w2[0] - warp[0]).abs().max().item() < 1e-4 and (c2[0] - cert[0]).abs().max().item() < 1e-4


@pytest.mark.gpu
def test_tiny_sample_gpu(golden, cuda):
    g = golden("tiny_e2e")
    model, _ = _build(cuda)
    A, B = _images(g)
    warp, cert = model.match(A.to(cuda), B.to(cuda), batched=False)
    m, c = model.sample(warp, cert, num=2000)
    assert m.shape == (2000, 4) and c.shape == (2000,)
    flat = warp.reshape(-1, 4)
    # every sampled row is a row of the dense warp
    d = (m[:50, None, :] - flat[None]).abs().sum(-1).min(dim=1).values
    assert d.max().item() == 0.0
    kA, kB = model.to_pixel_coordinates(m, 200, 232, 200, 232)
    assert kA.shape == (2000, 2) and kB.shape == (2000, 2)


@pytest.mark.gpu
def test_tiny_config5_batch_independence(cuda):
    """BASELINE config 5 shapes (480 x 640 pairs, coarse grid 60 x 80 = 4800 tokens per image): a batch of 8 pairs
    built from 2 distinct pairs reproduces the 2 single-pair results element for element (the fused correlation +
    soft-argmax kernel and the gathers treat batch elements independently), outputs finite and in range."""
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    model, _ = _build(cuda)
    g = torch.Generator().manual_seed(4321)
    A = torch.rand(2, 3, 480, 640, generator=g)
    B = torch.rand(2, 3, 480, 640, generator=g)
    w2, c2 = model.match(A.to(cuda), B.to(cuda), batched=True)
    assert w2.shape == (2, 480, 640, 4) and c2.shape == (2, 480, 640)
    assert torch.isfinite(w2).all() and torch.isfinite(c2).all()
    w8, c8 = model.match(A.repeat(4, 1, 1, 1).to(cuda), B.repeat(4, 1, 1, 1).to(cuda), batched=True)
    for i in range(8):
        assert (w8[i] - w2[i % 2]).abs().max().item() < 1e-4 and (c8[i] - c2[i % 2]).abs().max().item() < 1e-4
    m, c = model.sample(w8[3], c8[3], num=5000)
    assert m.shape == (5000, 4) and c.shape == (5000,)
