"""Tiny-RoMa (BASELINE config 5): host side around `roma_tiny_corr_posembed`,
`roma_refiner_gather` (warp of image-B features) and `roma_kde`.

Mirrors romatch/models/tiny.py:93-390 (`TinyRoMa`): same constructor arguments, state_dict keys
(`coarse_matcher.*`, `fine_matcher.*`), `match` / `sample` / `forward` behaviour.  The XFeat backbone is not part
of the reference tree (it is fetched with torch.hub from verlab/accelerated_features, model_zoo/__init__.py:22-26);
`XFeatBackbone` below restates the published XFeat feature trunk with the attribute names TinyRoMa reads
(`norm`, `skip1`, `block1..5`, `block_fusion`), so a real XFeat `.net` can be passed instead, and so that tests
can drive both this model and the reference's with the same seeded trunk.
"""
from __future__ import annotations

import math
from pathlib import Path
from typing import Optional

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F
from PIL import Image

from .. import ops


class BasicLayer(nn.Module):
    """Conv2d -> BatchNorm2d(affine=False) -> ReLU (tiny.py:78-91)."""

    def __init__(self, cin, cout, kernel_size=3, stride=1, padding=1, dilation=1, bias=False, relu=True):
        super().__init__()
        self.layer = nn.Sequential(
            nn.Conv2d(cin, cout, kernel_size, padding=padding, stride=stride, dilation=dilation, bias=bias),
            nn.BatchNorm2d(cout, affine=False),
            nn.ReLU(inplace=True) if relu else nn.Identity())

    def forward(self, x):
        return self.layer(x)


class XFeatBackbone(nn.Module):
    """XFeat feature trunk ("XFeat: Accelerated Features for Lightweight Image Matching", CVPR 2024), the part
    TinyRoMa.forward_single (tiny.py:144-161) evaluates: 1 -> 24 channels at stride 4 (x2) and 64 at stride 8."""

    def __init__(self):
        super().__init__()
        self.norm = nn.InstanceNorm2d(1)
        self.skip1 = nn.Sequential(nn.AvgPool2d(4, stride=4), nn.Conv2d(1, 24, 1, stride=1, padding=0))
        self.block1 = nn.Sequential(BasicLayer(1, 4, stride=1), BasicLayer(4, 8, stride=2),
                                    BasicLayer(8, 8, stride=1), BasicLayer(8, 24, stride=2))
        self.block2 = nn.Sequential(BasicLayer(24, 24, stride=1), BasicLayer(24, 24, stride=1))
        self.block3 = nn.Sequential(BasicLayer(24, 64, stride=2), BasicLayer(64, 64, stride=1),
                                    BasicLayer(64, 64, 1, padding=0))
        self.block4 = nn.Sequential(BasicLayer(64, 64, stride=2), BasicLayer(64, 64, stride=1),
                                    BasicLayer(64, 64, stride=1))
        self.block5 = nn.Sequential(BasicLayer(64, 128, stride=2), BasicLayer(128, 128, stride=1),
                                    BasicLayer(128, 128, stride=1), BasicLayer(128, 64, 1, padding=0))
        self.block_fusion = nn.Sequential(BasicLayer(64, 64, stride=1), BasicLayer(64, 64, stride=1),
                                          nn.Conv2d(64, 64, 1, padding=0))
        # heads the reference deletes on construction (tiny.py:105); kept as empty stubs so `del` works there too
        self.heatmap_head = nn.Identity()
        self.keypoint_head = nn.Identity()
        self.fine_matcher = nn.Identity()


def _grid(B, H, W, device):
    """Pixel-centre grid in [-1,1], (B,H,W,2) (tiny.py:278-284)."""
    xs = torch.linspace(-1 + 1 / W, 1 - 1 / W, W, device=device)
    ys = torch.linspace(-1 + 1 / H, 1 - 1 / H, H, device=device)
    return torch.stack(torch.meshgrid(xs, ys, indexing="xy"), dim=-1).float().expand(B, H, W, 2)


class TinyRoMa(nn.Module):
    """tiny.py:93-390.  `xfeat` is an XFeat `.net`-like module (default: a fresh `XFeatBackbone`)."""

    def __init__(self, xfeat: Optional[nn.Module] = None, freeze_xfeat=True, sample_mode="threshold_balanced",
                 symmetric=False, exact_softmax=False):
        super().__init__()
        xfeat = XFeatBackbone() if xfeat is None else xfeat
        for head in ("heatmap_head", "keypoint_head", "fine_matcher"):
            if hasattr(xfeat, head):
                delattr(xfeat, head)
        if freeze_xfeat:
            xfeat.train(False)
            self.xfeat = [xfeat]
        else:
            self.xfeat = nn.ModuleList([xfeat])
        self.freeze_xfeat = freeze_xfeat
        md, fd = 256, 64
        self.coarse_matcher = nn.Sequential(
            BasicLayer(64 + 64 + 2, md), BasicLayer(md, md), BasicLayer(md, md), BasicLayer(md, md),
            nn.Conv2d(md, 3, kernel_size=1, bias=True, padding=0))
        self.fine_matcher = nn.Sequential(
            BasicLayer(24 + 24 + 2, fd), BasicLayer(fd, fd), BasicLayer(fd, fd), BasicLayer(fd, fd),
            nn.Conv2d(fd, 3, kernel_size=1, bias=True, padding=0))
        self.sample_mode = sample_mode
        self.sample_thresh = 0.05
        self.symmetric = symmetric
        self.exact_softmax = exact_softmax

    @property
    def device(self):
        return self.fine_matcher[-1].weight.device

    def to(self, *args, **kwargs):
        if self.freeze_xfeat:
            self.xfeat[0].to(*args, **kwargs)     # hidden from .parameters(), moved explicitly
        return super().to(*args, **kwargs)

    def preprocess_tensor(self, x):
        """tiny.py:135-142: resize so both sides are multiples of 32."""
        H, W = x.shape[-2:]
        _H, _W = (H // 32) * 32, (W // 32) * 32
        return F.interpolate(x, (_H, _W), mode="bilinear", align_corners=False), H / _H, W / _W

    def forward_single(self, x):
        """tiny.py:144-161 -> (x2: (B,24,H/4,W/4), feats: (B,64,H/8,W/8))."""
        xf = self.xfeat[0]
        x = xf.norm(x.mean(dim=1, keepdim=True))
        x1 = xf.block1(x)
        x2 = xf.block2(x1 + xf.skip1(x))
        x3 = xf.block3(x2)
        x4 = xf.block4(x3)
        x5 = xf.block5(x4)
        x4 = F.interpolate(x4, x3.shape[-2:], mode="bilinear")
        x5 = F.interpolate(x5, x3.shape[-2:], mode="bilinear")
        return x2, xf.block_fusion(x3 + x4 + x5)

    # -- T1/T2 ----------------------------------------------------------------------------------------------
    def corr_volume(self, feat0, feat1):
        """tiny.py:241-254, materialised (exact_softmax / training only; inference uses the fused kernel)."""
        B, C, H0, W0 = feat0.shape
        _, _, H1, W1 = feat1.shape
        return torch.einsum("bci,bcj->bji", feat0.reshape(B, C, -1), feat1.reshape(B, C, -1)) \
            .reshape(B, H1, W1, H0, W0) / math.sqrt(C)

    def pos_embed(self, corr_volume):
        """tiny.py:178-199, exact branch (softmax over every image-1 position)."""
        B, H1, W1, H0, W0 = corr_volume.shape
        grid = _grid(1, H1, W1, corr_volume.device).reshape(H1 * W1, 2).to(corr_volume)
        P = corr_volume.reshape(B, H1 * W1, H0, W0).softmax(dim=1)
        return torch.einsum("bchw,cd->bdhw", P, grid)

    def coarse_warp(self, feat0, feat1):
        """pos_embed(corr_volume(...)): one fused kernel at inference (the volume is never formed)."""
        if self.exact_softmax or self.training:
            return self.pos_embed(self.corr_volume(feat0, feat1))
        return ops.tiny_corr_posembed(feat0, feat1)

    # -- forward / match --------------------------------------------------------------------------------------
    def forward(self, batch):
        """tiny.py:330-369 -> {8: {flow, certainty}, 4: {flow, certainty}}."""
        im0, _, _ = self.preprocess_tensor(batch["im_A"])
        im1, _, _ = self.preprocess_tensor(batch["im_B"])
        H1, W1 = im1.shape[-2:]
        to_normalized = torch.tensor((2 / W1, 2 / H1, 1), device=im0.device)[None, :, None, None]
        if im0.shape[-2:] == im1.shape[-2:]:
            f, c = self.forward_single(torch.cat([im0, im1], dim=0))
            (f0, f1), (c0, c1) = f.chunk(2), c.chunk(2)
        else:
            f0, c0 = self.forward_single(im0)
            f1, c1 = self.forward_single(im1)
        warp = self.coarse_warp(c0, c1)
        coarse = torch.cat((warp, torch.zeros_like(warp[:, -1:])), dim=1)
        c1w = ops.grid_sample(c1, coarse.permute(0, 2, 3, 1)[..., :2])
        coarse = coarse + self.coarse_matcher(torch.cat((c0, c1w, warp), dim=1)) * to_normalized
        out = {8: {"flow": coarse[:, :2], "certainty": coarse[:, 2:]}}
        up = F.interpolate(coarse, size=f0.shape[-2:], mode="bilinear", align_corners=False).detach()
        f1w = ops.grid_sample(f1, up.permute(0, 2, 3, 1)[..., :2])
        fine = up + self.fine_matcher(torch.cat((f0, f1w, up[:, :2]), dim=1)) * to_normalized
        out[4] = {"flow": fine[:, :2], "certainty": fine[:, 2:]}
        return out

    @torch.inference_mode()
    def match_from_path(self, im0_path, im1_path):
        from torchvision.transforms import ToTensor
        dev = self.device
        im0 = ToTensor()(Image.open(im0_path))[None].to(dev)
        im1 = ToTensor()(Image.open(im1_path))[None].to(dev)
        return self.match(im0, im1, batched=False)

    @torch.inference_mode()
    def match(self, im0, im1, *args, batched=True):
        """tiny.py:262-295: returns (warp (B,H0,W0,4), certainty (B,H0,W0)) from the stride-8 prediction,
        exactly as the reference does (the stride-4 refinement is computed but not used there either)."""
        if isinstance(im0, (str, Path)):
            return self.match_from_path(im0, im1)
        if isinstance(im0, Image.Image):
            from torchvision.transforms import ToTensor
            batched = False
            im0 = ToTensor()(im0)[None].to(self.device)
            im1 = ToTensor()(im1)[None].to(self.device)
        B, _, H0, W0 = im0.shape
        self.train(False)
        corresps = self.forward({"im_A": im0, "im_B": im1})
        flow = F.interpolate(corresps[8]["flow"], size=(H0, W0), mode="bilinear", align_corners=False) \
            .permute(0, 2, 3, 1).reshape(B, H0, W0, 2)
        cert = F.interpolate(corresps[8]["certainty"], size=(H0, W0), mode="bilinear", align_corners=False)
        warp, cert = torch.cat((_grid(B, H0, W0, flow.device), flow), dim=-1), cert[:, 0].sigmoid()
        return (warp, cert) if batched else (warp[0], cert[0])

    def sample(self, matches, certainty, num=5_000):
        """tiny.py:297-327; kde through `roma_kde` (half precision, down=1 on CUDA as the reference picks)."""
        if "threshold" in self.sample_mode:
            certainty = certainty.clone()
            certainty[certainty > self.sample_thresh] = 1
        matches, certainty = matches.reshape(-1, 4), certainty.reshape(-1)
        expansion = 4 if "balanced" in self.sample_mode else 1
        good = ops.multinomial_without_replacement(certainty, min(expansion * num, len(certainty)))
        good_matches, good_certainty = matches[good], certainty[good]
        if "balanced" not in self.sample_mode:
            return good_matches, good_certainty
        density = ops.kde(good_matches, std=0.1, half=True, down=1)
        p = 1 / (density + 1)
        p[density < 10] = 1e-7
        keep = ops.multinomial_without_replacement(p, min(num, len(good_certainty)))
        return good_matches[keep], good_certainty[keep]

    def to_pixel_coordinates(self, coords, H_A, W_A, H_B=None, W_B=None):
        if coords.shape[-1] == 2:
            return self._to_pixel_coordinates(coords, H_A, W_A)
        if isinstance(coords, (list, tuple)):
            kA, kB = coords[0], coords[1]
        else:
            kA, kB = coords[..., :2], coords[..., 2:]
        return self._to_pixel_coordinates(kA, H_A, W_A), self._to_pixel_coordinates(kB, H_B, W_B)

    @staticmethod
    def _to_pixel_coordinates(coords, H, W):
        return torch.stack((W / 2 * (coords[..., 0] + 1), H / 2 * (coords[..., 1] + 1)), dim=-1)


def random_init_tiny(seed: int = 0):
    """Seeded (TinyRoMa state_dict, XFeatBackbone state_dict) with non-trivial BatchNorm running statistics, so
    that the trunk does not collapse and the correlation softmax is not uniform."""
    g = torch.Generator().manual_seed(seed)
    model = TinyRoMa(XFeatBackbone())
    trunk = model.xfeat[0]
    for mod in (model, trunk):
        for name, p in list(mod.state_dict().items()):
            if name.endswith("num_batches_tracked"):
                continue
            if name.endswith("running_var"):
                p.copy_(0.5 + torch.rand(p.shape, generator=g))
            elif name.endswith("running_mean"):
                p.copy_(0.1 * torch.randn(p.shape, generator=g))
            elif p.dim() == 4:
                fan_in = p.shape[1] * p.shape[2] * p.shape[3]
                p.copy_(torch.randn(p.shape, generator=g) * math.sqrt(2.0 / fan_in))
            else:
                p.copy_(0.05 * torch.randn(p.shape, generator=g))
    return {k: v.clone() for k, v in model.state_dict().items()}, {k: v.clone() for k, v in trunk.state_dict().items()}
