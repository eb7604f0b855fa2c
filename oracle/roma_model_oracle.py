"""CPU oracle of the WHOLE matching path — TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Only `tests/`, `__graft_entry__.smoke()` and the `cpu_baseline` / `--impl reference` legs of `bench.py` may
import this module.  Nothing under `roma_b200/` imports it.

It restates `RegressionMatcher.match` / `Decoder.forward` / `ConvRefiner.forward` of the reference on the CPU
in fp32 — the reference's own CPU configuration (model_zoo/__init__.py:36-37 forces fp32, utils.py:640-654
switches autocast off) — composing
  * the hot-path restatements of `oracle/roma_oracle.py` (GP, grid_sample, local_correlation, update /
    resample, epilogue, cls_to_flow_refine), and
  * the PyTorch bodies that the product keeps as PyTorch (VGG19 / DINOv2 encoder, proj, transformer decoder,
    refiner conv stacks) taken from a `roma_b200` model instance, executed on the CPU.
It never calls a CUDA kernel and never calls `model.match` / `model.decoder.forward` of the product.

Pinned by `tests/golden/make_golden_e2e.py`, which runs the unmodified reference (`/root/reference`) with the
same seeded weights in the build container and commits its outputs as `tests/golden/e2e_*.npz`.
"""
from __future__ import annotations

import math
from typing import Dict, Optional

import torch

from . import roma_oracle as O


def refiner_forward(ref, x, y, flow, scale_factor=1.0):
    """ConvRefiner.forward — matcher.py:105-143 (CPU: autocast off, everything fp32)."""
    dw = ref.disp_emb.weight if ref.has_displacement_emb else None
    db = ref.disp_emb.bias if ref.has_displacement_emb else None
    d = O.refiner_inputs(x.float(), y.float(), flow, dw, db, 40 / 32 * scale_factor, ref.local_corr_radius)
    d = ref.hidden_blocks(ref.block1(d))
    d = ref.out_conv(d.float())
    return d[:, :-1], d[:, -1:]


def embedding_decoder_forward(dec, gp_posterior, features):
    """TransformerDecoder.forward — transformer/__init__.py:30-46."""
    B, _, H, W = gp_posterior.shape
    x = torch.cat((gp_posterior, features), dim=1)
    tokens = x.reshape(B, x.shape[1], H * W).permute(0, 2, 1)
    out = dec.to_out(dec.blocks(tokens))
    out = out.permute(0, 2, 1).reshape(B, dec.out_dim, H, W)
    return out[:, :-1], out[:, -1:]


def decoder_forward(decoder, f1: Dict[int, torch.Tensor], f2: Dict[int, torch.Tensor], upsample=False,
                    flow=None, certainty=None, scale_factor=1.0):
    """Decoder.forward — matcher.py:333-422."""
    all_scales = decoder.scales if not upsample else ["8", "4", "2", "1"]
    sizes = {s: f1[s].shape[-2:] for s in f1}
    h, w = sizes[1]
    corresps = {}
    if upsample:   # :349-360
        flow = O.bilinear_resize(flow, sizes[int(all_scales[0])])
        certainty = O.bilinear_resize(certainty, sizes[int(all_scales[0])])
    else:
        certainty = 0.0
    for new_scale in all_scales:
        ins = int(new_scale)
        f1_s, f2_s = f1[ins].float(), f2[ins].float()
        if new_scale in decoder.proj:   # :366-371
            f1_s, f2_s = decoder.proj[new_scale](f1_s), decoder.proj[new_scale](f2_s)
        if ins == 16:   # :373-389
            gp = decoder.gps[new_scale]
            mu = O.gp_posterior(f1_s, f2_s, gp.pos_conv.weight, gp.pos_conv.bias, T=gp.K.T, sigma_noise=gp.sigma_noise)
            cls, certainty = embedding_decoder_forward(decoder.embedding_decoder, mu, f1_s)
            flow = O.cls_to_flow_refine(cls).permute(0, 3, 1, 2)
        if new_scale in decoder.conv_refiner:   # :391-402
            dflow, dcert = refiner_forward(decoder.conv_refiner[new_scale], f1_s, f2_s, flow, scale_factor)
            flow, certainty = O.flow_update(flow, certainty, torch.cat((dflow, dcert), dim=1), ins, h, w)
        corresps[ins] = {"certainty": certainty, "flow": flow}
        if new_scale != "1":   # :407-417
            flow = O.bilinear_resize(flow, sizes[ins // 2])
            certainty = O.bilinear_resize(certainty, sizes[ins // 2])
    return corresps


def encoder_forward(encoder, x, upsample=False):
    """CNNandDinov2.forward — encoders.py:110-121 (the bodies are plain PyTorch; run as they are on the CPU)."""
    return {s: f.float() for s, f in encoder(x, upsample=upsample).items()}


def forward_symmetric(model, im_A, im_B, upsample=False, scale_factor=1.0, corresps=None):
    """RegressionMatcher.forward_symmetric — matcher.py:516-528."""
    pyr = encoder_forward(model.encoder, torch.cat((im_A, im_B), dim=0), upsample)
    f_s = {s: torch.cat((f.chunk(2)[1], f.chunk(2)[0]), dim=0) for s, f in pyr.items()}
    return decoder_forward(model.decoder, pyr, f_s, upsample=upsample, scale_factor=scale_factor, **(corresps or {}))


def forward_asymmetric(model, im_A, im_B, upsample=False, scale_factor=1.0, corresps=None):
    """RegressionMatcher.forward — matcher.py:496-514."""
    pyr = encoder_forward(model.encoder, torch.cat((im_A, im_B), dim=0), upsample)
    f_q = {s: f.chunk(2)[0] for s, f in pyr.items()}
    f_s = {s: f.chunk(2)[1] for s, f in pyr.items()}
    return decoder_forward(model.decoder, f_q, f_s, upsample=upsample, scale_factor=scale_factor, **(corresps or {}))


@torch.inference_mode()
def match_tensors(model, coarse_A, coarse_B, fine_A: Optional[torch.Tensor] = None,
                  fine_B: Optional[torch.Tensor] = None, return_logits: bool = False):
    """RegressionMatcher.match — matcher.py:593-730 — from already normalised CPU tensors:
    coarse_* (B,3,h_resized,w_resized); fine_* (B,3,*upsample_res) when model.upsample_preds.
    Returns (warp (B,H,Wout,4), certainty (B,H,Wout)[, logits])."""
    fwd = forward_symmetric if model.symmetric else forward_asymmetric
    corresps = fwd(model, coarse_A.float(), coarse_B.float())
    cert16 = corresps[16]["certainty"] if model.attenuate_cert else None
    if model.upsample_preds:
        hs, ws = model.upsample_res
        sf = math.sqrt(hs * ws / (model.w_resized * model.h_resized))
        corresps = fwd(model, fine_A.float(), fine_B.float(), upsample=True, scale_factor=sf, corresps=corresps[1])
    warp, cert, logits = O.match_epilogue(corresps[1]["flow"], corresps[1]["certainty"], cert16, model.symmetric)
    return (warp, cert, logits) if return_logits else (warp, cert)


# ----------------------------------------------------------------------------------------------
# Tiny-RoMa (BASELINE config 5)
# ----------------------------------------------------------------------------------------------
def tiny_forward(model, im0: torch.Tensor, im1: torch.Tensor):
    """TinyRoMa.forward — romatch/models/tiny.py:330-369 (inference branch of pos_embed, :186-197).
    `model` is a roma_b200.host.tiny.TinyRoMa on the CPU; only its torch modules are used."""
    import torch.nn.functional as F
    im0, _, _ = model.preprocess_tensor(im0)
    im1, _, _ = model.preprocess_tensor(im1)
    H1, W1 = im1.shape[-2:]
    to_normalized = torch.tensor((2 / W1, 2 / H1, 1.0))[None, :, None, None]
    f0, c0 = model.forward_single(im0)
    f1, c1 = model.forward_single(im1)
    warp = O.tiny_pos_embed(O.tiny_corr_volume(c0, c1))
    coarse = torch.cat((warp, torch.zeros_like(warp[:, -1:])), dim=1)
    c1w = O.grid_sample_zeros(c1, coarse.permute(0, 2, 3, 1)[..., :2])
    coarse = coarse + model.coarse_matcher(torch.cat((c0, c1w, warp), dim=1)) * to_normalized
    out = {8: {"flow": coarse[:, :2], "certainty": coarse[:, 2:]}}
    up = F.interpolate(coarse, size=f0.shape[-2:], mode="bilinear", align_corners=False)
    f1w = O.grid_sample_zeros(f1, up.permute(0, 2, 3, 1)[..., :2])
    fine = up + model.fine_matcher(torch.cat((f0, f1w, up[:, :2]), dim=1)) * to_normalized
    out[4] = {"flow": fine[:, :2], "certainty": fine[:, 2:]}
    return out


def tiny_match(model, im0: torch.Tensor, im1: torch.Tensor):
    """TinyRoMa.match — tiny.py:262-295 -> (warp (B,H0,W0,4), certainty (B,H0,W0))."""
    B, _, H0, W0 = im0.shape
    c = tiny_forward(model, im0, im1)[8]
    flow = O.bilinear_resize(c["flow"], (H0, W0)).permute(0, 2, 3, 1)
    cert = O.bilinear_resize(c["certainty"], (H0, W0))[:, 0].sigmoid()
    xs, ys = O.pixel_centers(W0, flow.device), O.pixel_centers(H0, flow.device)
    grid = torch.stack((xs[None, :].expand(H0, W0), ys[:, None].expand(H0, W0)), dim=-1)[None].expand(B, H0, W0, 2)
    return torch.cat((grid, flow), dim=-1), cert
