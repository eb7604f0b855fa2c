"""GPU: the CUDA product path (roma_b200 model -> C ABI kernels) against the CPU full-path oracle and the
reference's own outputs (tests/golden/e2e_*.npz), on identical seeded weights and inputs.

Tolerances are north_star's: warp <= 1e-3 in normalised coordinates, certainty logits <= 1e-2.  All PyTorch
bodies run in fp32 here (amp_dtype=float32) so the comparison isolates the kernels; bf16/fp16 feature inputs are
covered kernel by kernel in test_kernels_gpu.py."""
import numpy as np
import pytest
import torch
from PIL import Image

import roma_b200
from roma_b200.matcher import _pil_to_tensor
from oracle import roma_model_oracle as MO

pytestmark = pytest.mark.gpu
WARP_TOL, LOGIT_TOL = 1e-3, 1e-2


@pytest.fixture(scope="module", params=["default", "sensitised"])
def case(request, golden, cuda):
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    g = golden(f"e2e_{request.param}")
    weights, dinov2 = roma_b200.random_init_weights(seed=0, sensitised=request.param == "sensitised")
    assert roma_b200.state_digest(weights) == bytes(g["digest"].numpy().astype(np.uint8)).decode()
    coarse, up = tuple(int(v) for v in g["coarse"]), tuple(int(v) for v in g["upsample"])
    cpu = roma_b200.roma_outdoor("cpu", weights, dinov2, coarse_res=coarse, upsample_res=up, amp_dtype=torch.float32)
    gpu = roma_b200.roma_outdoor(cuda, weights, dinov2, coarse_res=coarse, upsample_res=up, amp_dtype=torch.float32)
    return g, cpu, gpu, coarse, up


def test_match_paths_vs_reference_golden(case, tmp_path):
    """model.match(pathA, pathB) through PIL decoding, exactly the reference call (matcher.py:593-730)."""
    g, _, gpu, _, _ = case
    Image.fromarray(g["im_A"].numpy()).save(tmp_path / "A.png")
    Image.fromarray(g["im_B"].numpy()).save(tmp_path / "B.png")
    warp, cert = gpu.match(str(tmp_path / "A.png"), str(tmp_path / "B.png"), device="cuda")
    assert warp.shape == (168, 392, 4) and cert.shape == (168, 392)
    assert (warp[::2, ::2].cpu() - g["warp"]).abs().max().item() < WARP_TOL
    assert (cert[::2, ::2].cpu() - g["certainty"]).abs().max().item() < LOGIT_TOL   # probabilities: tighter than logits


def test_match_paths_device_resize_is_pil_exact(case, tmp_path):
    """File inputs: PIL's resize + ToTensor + Normalize on the device (roma_prepare_images_pil) feed the encoder the very
    same pixels as the host PIL path, so the whole match is bit-identical between the two."""
    from roma_b200 import ops
    g, _, gpu, coarse, up = case
    Image.fromarray(g["im_A"].numpy()).save(tmp_path / "A.png")
    Image.fromarray(g["im_B"].numpy()).save(tmp_path / "B.png")
    assert gpu.native_image_prep
    for name in ("A.png", "B.png"):                      # the encoder inputs themselves: bit for bit
        pil = Image.open(tmp_path / name).convert("RGB")
        u8 = torch.from_numpy(np.array(pil)).cuda()
        for res in (coarse, up):
            host = gpu._normalise(_pil_to_tensor(pil, tuple(res))[None].cuda())
            assert torch.equal(ops.prepare_images_pil(u8, tuple(res), channels_last=True), host)
    warp, cert = gpu.match(str(tmp_path / "A.png"), str(tmp_path / "B.png"), device="cuda")
    gpu.native_image_prep = False
    try:
        warp_h, cert_h = gpu.match(str(tmp_path / "A.png"), str(tmp_path / "B.png"), device="cuda")
    finally:
        gpu.native_image_prep = True
    # identical inputs through the same kernels: equal up to the run-to-run reproducibility of the library convolutions
    assert (warp - warp_h).abs().max().item() <= 1e-6 and (cert - cert_h).abs().max().item() <= 1e-6


def test_match_tensors_vs_oracle_logits(case):
    g, cpu, gpu, coarse, up = case
    prep = lambda im, res: cpu._normalise(_pil_to_tensor(Image.fromarray(im.numpy()), tuple(res))[None])
    cA, cB, fA, fB = prep(g["im_A"], coarse), prep(g["im_B"], coarse), prep(g["im_A"], up), prep(g["im_B"], up)
    warp_o, cert_o, logit_o = MO.match_tensors(cpu, cA, cB, fA, fB, return_logits=True)
    # batched tensor input at the upsample resolution; the coarse input is what the reference would have produced
    gpu_coarse = {"im_A": cA.cuda(), "im_B": cB.cuda()}
    corr = gpu.forward_symmetric(gpu_coarse)
    corr_o = MO.forward_symmetric(cpu, cA, cB)
    for s in (16, 8, 4, 2, 1):
        assert (corr[s]["flow"].cpu() - corr_o[s]["flow"]).abs().max().item() < WARP_TOL, f"scale {s} flow"
        assert (corr[s]["certainty"].cpu() - corr_o[s]["certainty"]).abs().max().item() < LOGIT_TOL, f"scale {s} certainty"


def test_single_pass_asymmetric(case, tmp_path):
    g, _, gpu, _, _ = case
    Image.fromarray(g["im_A"].numpy()).save(tmp_path / "A.png")
    Image.fromarray(g["im_B"].numpy()).save(tmp_path / "B.png")
    gpu.symmetric, gpu.upsample_preds, gpu.attenuate_cert = False, False, False
    try:
        warp, cert = gpu.match(str(tmp_path / "A.png"), str(tmp_path / "B.png"), device="cuda")
    finally:
        gpu.symmetric, gpu.upsample_preds, gpu.attenuate_cert = True, True, True
    assert (warp[::2, ::2].cpu() - g["warp_single"]).abs().max().item() < WARP_TOL
    assert (cert[::2, ::2].cpu() - g["certainty_single"]).abs().max().item() < LOGIT_TOL


def test_batched_uint8_tensor_inputs(case):
    """Tensor / batched inputs (broken in the reference): a batch of 2 equals two single calls."""
    g, _, gpu, _, up = case
    t = lambda im: torch.from_numpy(np.asarray(Image.fromarray(im.numpy()).resize((up[1], up[0]), Image.BICUBIC))).permute(2, 0, 1)
    A, B = t(g["im_A"]), t(g["im_B"])
    bA, bB = torch.stack((A, B)), torch.stack((B, A))
    warp, cert = gpu.match(bA, bB, batched=True, device="cuda")
    w0, c0 = gpu.match(A[None], B[None], batched=True, device="cuda")
    assert warp.shape == (2, up[0], 2 * up[1], 4)
    assert (warp[0] - w0[0]).abs().max().item() < 1e-4 and (cert[0] - c0[0]).abs().max().item() < 1e-3


def test_visualize_warp_and_fb_consistency_headless(tmp_path):
    """§8(f)-4: `visualize_warp` (matcher.py:732-766) and `conf_from_fb_consistency` (:530-551) run on the K2 gather
    kernel; checked against F.grid_sample on the CPU.  Headless: the visualisation is a tensor, optionally a file."""
    import torch.nn.functional as F
    from roma_b200.matcher import RegressionMatcher
    dev = torch.device("cuda")
    m = RegressionMatcher(torch.nn.Identity(), torch.nn.Identity(), h=24, w=32, symmetric=True)
    g = torch.Generator().manual_seed(5)
    H, W = 24, 32
    warp = torch.rand(H, 2 * W, 4, generator=g) * 2.2 - 1.1
    cert = torch.rand(H, 2 * W, generator=g)
    xA, xB = torch.rand(3, H, W, generator=g), torch.rand(3, H, W, generator=g)
    vis = m.visualize_warp(warp.to(dev), cert.to(dev), im_A=xA.to(dev), im_B=xB.to(dev), symmetric=True,
                           save_path=str(tmp_path / "vis.png"))
    a = F.grid_sample(xB[None], warp[None, :, :W, 2:], mode="bilinear", align_corners=False)[0]
    b = F.grid_sample(xA[None], warp[None, :, W:, :2], mode="bilinear", align_corners=False)[0]
    want = cert * torch.cat((a, b), dim=2) + (1 - cert)
    assert vis.shape == (3, H, 2 * W) and (vis.cpu() - want).abs().max().item() < 1e-5
    assert (tmp_path / "vis.png").exists()
    vis1 = m.visualize_warp(warp[:, :W].to(dev), cert[:, :W].to(dev), im_A=xA.to(dev), im_B=xB.to(dev), symmetric=False)
    assert (vis1.cpu() - (cert[:, :W] * a + (1 - cert[:, :W]))).abs().max().item() < 1e-5

    fwd = torch.rand(2, H, W, 2, generator=g) * 2 - 1
    bwd = torch.rand(2, H, W, 2, generator=g) * 2 - 1
    got = m.conf_from_fb_consistency(fwd.to(dev), bwd.to(dev), th=6)
    xs = torch.linspace(-1 + 1 / W, 1 - 1 / W, W); ys = torch.linspace(-1 + 1 / H, 1 - 1 / H, H)
    coords = torch.stack((xs[None, :].expand(H, W), ys[:, None].expand(H, W)), dim=-1)
    fb = F.grid_sample(bwd.permute(0, 3, 1, 2), fwd, mode="bilinear", align_corners=False).permute(0, 2, 3, 1)
    dist = (coords - fb).norm(dim=-1)
    want_mask = (dist < 2 * 6 / max(H, W)).float()
    sure = (dist - 2 * 6 / max(H, W)).abs() > 1e-5                  # away from the threshold
    assert got.shape == (2, H, W) and torch.equal(got.cpu()[sure], want_mask[sure])
