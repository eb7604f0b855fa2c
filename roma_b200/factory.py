"""Model factories with the reference's names and keyword arguments
(romatch/models/model_zoo/__init__.py:16-73, roma_models.py:9-170).

`weights` / `dinov2_weights` are the same state_dicts the reference takes.  Offline there is nothing to
download, so `random_init_weights()` builds seeded random-init state_dicts with the reference's key set
(603 tensors for roma_outdoor; the DINOv2 dict separately), which is what BASELINE.json's configs use.
"""
from __future__ import annotations

from typing import Dict, Optional, Tuple, Union

import torch
import torch.nn as nn

from .host.decoder import build_decoder
from .host.encoder import CNNandDinov2
from .host.vit import DinoViT
from .host.tiny import TinyRoMa, XFeatBackbone
from .matcher import RegressionMatcher

weight_urls = {
    "romatch": {
        "outdoor": "https://github.com/Parskatt/storage/releases/download/roma/roma_outdoor.pth",
        "indoor": "https://github.com/Parskatt/storage/releases/download/roma/roma_indoor.pth",
    },
    "tiny_roma_v1": {
        "outdoor": "https://github.com/Parskatt/storage/releases/download/roma/tiny_roma_v1_outdoor.pth",
    },
    "dinov2": "https://dl.fbaipublicfiles.com/dinov2/dinov2_vitl14/dinov2_vitl14_pretrain.pth",
}

Res = Union[int, Tuple[int, int]]


def _pair(r: Res) -> Tuple[int, int]:
    return (r, r) if isinstance(r, int) else tuple(r)


def roma_model(resolution, upsample_preds, device=None, weights=None, dinov2_weights=None,
               amp_dtype: torch.dtype = torch.float16, **kwargs) -> RegressionMatcher:
    """roma_models.py:18-170.  Unlike the reference (which hard-wires fp16 into the decoder / refiners / VGG,
    SURVEY.md §0), `amp_dtype` applies to every PyTorch body, so bf16 means bf16 everywhere."""
    decoder = build_decoder(amp_dtype)
    encoder = CNNandDinov2(dinov2_weights=dinov2_weights, amp_dtype=amp_dtype)
    h, w = resolution
    matcher = RegressionMatcher(encoder, decoder, h=h, w=w, upsample_preds=upsample_preds, symmetric=True,
                                attenuate_cert=True, sample_mode="threshold_balanced", **kwargs)
    if weights is not None:
        matcher.load_state_dict(weights)
    matcher = matcher.to(device)
    if device is not None and torch.device(device).type == "cuda":
        matcher = matcher.to(memory_format=torch.channels_last)
    return matcher.eval()


def _roma(kind: str, device, weights, dinov2_weights, coarse_res: Res, upsample_res: Res, amp_dtype):
    coarse_res, upsample_res = _pair(coarse_res), _pair(upsample_res)
    assert coarse_res[0] % 14 == 0 and coarse_res[1] % 14 == 0, "Needs to be multiple of 14 for backbone"
    if weights is None:
        weights = torch.hub.load_state_dict_from_url(weight_urls["romatch"][kind], map_location=device)
    if dinov2_weights is None:
        dinov2_weights = torch.hub.load_state_dict_from_url(weight_urls["dinov2"], map_location=device)
    model = roma_model(resolution=coarse_res, upsample_preds=True, weights=weights, dinov2_weights=dinov2_weights,
                       device=device, amp_dtype=amp_dtype)
    model.upsample_res = upsample_res
    return model


def roma_outdoor(device, weights=None, dinov2_weights=None, coarse_res: Res = 560, upsample_res: Res = 864,
                 amp_dtype: torch.dtype = torch.float16) -> RegressionMatcher:
    """model_zoo/__init__.py:30-52."""
    return _roma("outdoor", device, weights, dinov2_weights, coarse_res, upsample_res, amp_dtype)


def roma_indoor(device, weights=None, dinov2_weights=None, coarse_res: Res = 560, upsample_res: Res = 864,
                amp_dtype: torch.dtype = torch.float16) -> RegressionMatcher:
    """model_zoo/__init__.py:54-73."""
    return _roma("indoor", device, weights, dinov2_weights, coarse_res, upsample_res, amp_dtype)


# ------------------------------------------------------------------------------------------------------
# seeded random-init weights (BASELINE.json configs are all "random-init": no checkpoints offline)
# ------------------------------------------------------------------------------------------------------
def random_init_weights(seed: int = 0, sensitised: bool = False) -> Tuple[Dict[str, torch.Tensor], Dict[str, torch.Tensor]]:
    """(weights, dinov2_weights) for roma_outdoor, deterministic in `seed` (CPU generator, construction order).

    sensitised=False is PyTorch's default init — under which the final warp is blind to the refiner inputs
    (SURVEY.md §7.4-3).  sensitised=True re-draws every Conv2d of decoder.conv_refiner / decoder.proj He-normal
    with zero bias and scales out_conv by 0.01 so that the refiner gather kernels influence the output."""
    torch.manual_seed(seed)
    vit = DinoViT()
    dinov2 = {k: v.detach().clone() for k, v in vit.state_dict().items()}
    model = roma_model((560, 560), True, device=None, weights=None, dinov2_weights=vit, amp_dtype=torch.float32)
    if sensitised:
        g = torch.Generator().manual_seed(seed + 1)
        for name, m in model.decoder.named_modules():
            if isinstance(m, nn.Conv2d) and (name.startswith("conv_refiner") or name.startswith("proj")):
                fan_in = m.weight.shape[1] * m.weight.shape[2] * m.weight.shape[3]
                with torch.no_grad():
                    m.weight.copy_(torch.randn(m.weight.shape, generator=g) * (2.0 / fan_in) ** 0.5)
                    if m.bias is not None:
                        m.bias.zero_()
                    if name.endswith("out_conv"):
                        m.weight.mul_(0.01)
    weights = {k: v.detach().clone() for k, v in model.state_dict().items()}
    return weights, dinov2


def state_digest(sd: Dict[str, torch.Tensor]) -> str:
    """Order-independent fingerprint of a state_dict (used by the golden fixtures to prove identical weights)."""
    import hashlib
    h = hashlib.sha256()
    for k in sorted(sd):
        h.update(k.encode())
        h.update(sd[k].detach().cpu().contiguous().numpy().tobytes())
    return h.hexdigest()[:16]


def tiny_roma_v1_model(weights=None, freeze_xfeat=False, exact_softmax=False, xfeat=None):
    """roma_models.py:9-16."""
    model = TinyRoMa(xfeat=xfeat, freeze_xfeat=freeze_xfeat, exact_softmax=exact_softmax)
    if weights is not None:
        model.load_state_dict(weights)
    return model


def tiny_roma_v1_outdoor(device, weights=None, xfeat=None):
    """model_zoo/__init__.py:16-28.  Offline: `weights` is the reference's `tiny_roma_v1_outdoor.pth` state_dict
    and `xfeat` an XFeat `.net` (or an `XFeatBackbone` with XFeat's weights loaded); with neither, a seeded
    random init of the same architecture (`roma_b200.host.tiny.random_init_tiny`) is used."""
    if weights is None and xfeat is None:
        from .host.tiny import random_init_tiny
        weights, trunk_sd = random_init_tiny(seed=0)
        xfeat = XFeatBackbone()
        weights = {**weights, **{"xfeat.0." + k: v for k, v in trunk_sd.items()}}   # trunk is trainable here (:9)
    elif xfeat is None:
        raise ValueError("tiny_roma_v1_outdoor: pass the XFeat trunk (xfeat=...) together with trained weights; "
                         "torch.hub is not reachable from this build")
    return tiny_roma_v1_model(weights=weights, xfeat=xfeat).to(device)
