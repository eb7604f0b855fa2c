"""CPU: the full-path oracle (oracle/roma_model_oracle.py) against the reference's own `match()` outputs
(tests/golden/e2e_*.npz, produced by tests/golden/make_golden_e2e.py from /root/reference)."""
import numpy as np
import pytest
import torch
from PIL import Image

import roma_b200
from roma_b200.matcher import _pil_to_tensor
from oracle import roma_model_oracle as MO

WARP_TOL, CERT_TOL = 1e-3, 1e-2      # north_star tolerances (normalised coords / logits); the oracle is far inside


def prepare(model, im, res):
    return model._normalise(_pil_to_tensor(Image.fromarray(im.numpy()), tuple(res))[None])


@pytest.fixture(scope="module", params=["default", "sensitised"])
def case(request, golden):
    g = golden(f"e2e_{request.param}")
    torch.set_num_threads(8)
    weights, dinov2 = roma_b200.random_init_weights(seed=0, sensitised=request.param == "sensitised")
    want = bytes(g["digest"].numpy().astype(np.uint8)).decode()
    assert roma_b200.state_digest(weights) == want, "seeded init is not reproducible on this torch build"
    coarse, up = tuple(int(v) for v in g["coarse"]), tuple(int(v) for v in g["upsample"])
    model = roma_b200.roma_outdoor("cpu", weights, dinov2, coarse_res=coarse, upsample_res=up, amp_dtype=torch.float32)
    return g, model, coarse, up


def test_oracle_match_equals_reference(case):
    g, model, coarse, up = case
    cA, cB = prepare(model, g["im_A"], coarse), prepare(model, g["im_B"], coarse)
    fA, fB = prepare(model, g["im_A"], up), prepare(model, g["im_B"], up)
    warp, cert = MO.match_tensors(model, cA, cB, fA, fB)
    assert (warp[0, ::2, ::2] - g["warp"]).abs().max().item() < 1e-4
    assert (cert[0, ::2, ::2] - g["certainty"]).abs().max().item() < 1e-4
    # other branches of match(): asymmetric, single pass, no attenuation (matcher.py:653, 664, 719-720)
    model.symmetric, model.upsample_preds, model.attenuate_cert = False, False, False
    warp1, cert1 = MO.match_tensors(model, cA, cB)
    assert (warp1[0, ::2, ::2] - g["warp_single"]).abs().max().item() < 1e-4
    assert (cert1[0, ::2, ::2] - g["certainty_single"]).abs().max().item() < 1e-4


def test_product_has_no_cpu_path(case):
    _, model, _, _ = case
    with pytest.raises(RuntimeError):
        model.match(torch.zeros(1, 3, 112, 126), torch.zeros(1, 3, 112, 126), device="cpu")
