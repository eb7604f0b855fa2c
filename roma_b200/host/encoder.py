"""Feature pyramid encoder (stays PyTorch per north_star): VGG19-bn slices at strides 1/2/4/8 plus DINOv2
ViT-L/14 patch tokens as the "scale 16" map (really stride 14).

Mirrors romatch/models/encoders.py:61-121 (state_dict keys `cnn.layers.N.*`; the ViT is deliberately NOT a
registered sub-module there — encoders.py:104 — so its weights travel separately as `dinov2_weights`).
Differences that do not change results: the last max-pool of features[:40] (whose output the reference
discards) is not executed, and feature maps are produced channels-last because the kernels consume NHWC.
"""
from __future__ import annotations

from typing import Dict, Optional

import torch
import torch.nn as nn
import torch.nn.functional as F

from .. import ops
import torchvision.models as tvm

from .vit import DinoViT


def _autocast(device: torch.device, dtype: torch.dtype):
    """The reference forces autocast on for CUDA and off elsewhere (utils.py:640-654); fp32 means 'off'."""
    on = device.type == "cuda" and dtype in (torch.float16, torch.bfloat16)
    return torch.autocast(device.type if device.type == "cuda" else "cpu", enabled=on,
                          dtype=dtype if on else torch.bfloat16)


class VGG19(nn.Module):
    """VGG19-bn trunk (encoders.py:61-78).  On CUDA at inference the conv -> BatchNorm(eval) -> ReLU triples run from
    a prepared copy of the weights: BatchNorm folded into the conv, bias + ReLU fused into cuDNN's conv epilogue
    (one library launch per layer instead of four).  Same function; the parameters stay the reference's."""

    def __init__(self, amp_dtype: torch.dtype = torch.float16):
        super().__init__()
        self.layers = nn.ModuleList(tvm.vgg19_bn(weights=None).features[:40])
        self.amp_dtype = amp_dtype
        self._prep, self._prep_key = None, None

    def _prepare(self, device, dtype):
        key = (device, dtype, tuple(t._version for t in self.state_dict(keep_vars=True).values()),
               tuple(t.data_ptr() for t in self.parameters()))
        if self._prep_key == key:
            return self._prep
        plan, i = [], 0
        with torch.inference_mode(False), torch.no_grad():   # cached tensors must be normal (not inference) tensors
            while i < len(self.layers):
                layer = self.layers[i]
                if isinstance(layer, nn.MaxPool2d):
                    plan.append(("pool", layer))
                    i += 1
                    continue
                conv, bn = layer, self.layers[i + 1]
                assert isinstance(conv, nn.Conv2d) and isinstance(bn, nn.BatchNorm2d) and isinstance(self.layers[i + 2], nn.ReLU)
                scale = bn.weight.float() / torch.sqrt(bn.running_var.float() + bn.eps)
                w = conv.weight.float() * scale.view(-1, 1, 1, 1)
                if w.shape[1] % 8 and dtype != torch.float32:
                    # RGB stem: 3 input channels keep cuDNN on a legacy (sm80) kernel; zero-padding the input
                    # channels to 8 puts it on the tensor-core implicit GEMM.  The input is padded to match below.
                    w = F.pad(w, (0, 0, 0, 0, 0, 8 - w.shape[1] % 8))
                w = w.to(device=device, dtype=dtype).contiguous(memory_format=torch.channels_last)
                b = ((conv.bias.float() if conv.bias is not None else 0) * scale + bn.bias.float() - bn.running_mean.float() * scale)
                plan.append(("conv", w, b.to(device=device, dtype=dtype), conv.padding))
                i += 3
        self._prep, self._prep_key = plan, key
        return plan

    def _forward_prepared(self, x: torch.Tensor) -> Dict[int, torch.Tensor]:
        dtype = self.amp_dtype if self.amp_dtype in (torch.float16, torch.bfloat16) else torch.float32
        plan = self._prepare(x.device, dtype)
        x = x.to(dtype).contiguous(memory_format=torch.channels_last)
        feats, scale = {}, 1
        for n, step in enumerate(plan):
            if step[0] == "pool":
                feats[scale] = x
                scale *= 2
                if n == len(plan) - 1:
                    break
                x = ops.maxpool2x2(x) if x.shape[1] % 8 == 0 else step[1](x)
            else:
                if step[1].shape[1] != x.shape[1]:
                    x = F.pad(x, (0, 0, 0, 0, 0, step[1].shape[1] - x.shape[1])).contiguous(memory_format=torch.channels_last)
                x = torch.cudnn_convolution_relu(x, step[1], step[2], (1, 1), step[3], (1, 1), 1)
        return feats

    def forward(self, x: torch.Tensor) -> Dict[int, torch.Tensor]:
        if x.is_cuda and not self.training:
            return self._forward_prepared(x)
        feats, scale = {}, 1
        last = len(self.layers) - 1
        with _autocast(x.device, self.amp_dtype):
            for i, layer in enumerate(self.layers):
                if isinstance(layer, nn.MaxPool2d):
                    feats[scale] = x
                    scale *= 2
                    if i == last:
                        break
                x = layer(x)
        return feats


class CNNandDinov2(nn.Module):
    def __init__(self, dinov2_weights: Optional[dict] = None, amp_dtype: torch.dtype = torch.float16):
        super().__init__()
        if isinstance(dinov2_weights, DinoViT):
            vit = dinov2_weights.eval()
        else:
            vit = DinoViT(init=dinov2_weights is None).eval()
            if dinov2_weights is not None:
                vit.load_state_dict(dinov2_weights)
        self.cnn = VGG19(amp_dtype)
        self.amp_dtype = amp_dtype
        self.dinov2_vitl14 = [vit]          # hidden from state_dict / DDP like encoders.py:104

    def train(self, mode: bool = True):
        self.cnn.train(mode)
        return self

    def _vit(self, device) -> DinoViT:
        vit = self.dinov2_vitl14[0]
        want = self.amp_dtype if device.type == "cuda" else torch.float32
        p = vit.cls_token
        if p.device != device or p.dtype != want:
            vit = vit.to(device=device, dtype=want)
            self.dinov2_vitl14[0] = vit
        return vit

    def forward(self, x: torch.Tensor, upsample: bool = False) -> Dict[int, torch.Tensor]:
        B, _, H, W = x.shape
        if x.is_cuda:
            x = x.contiguous(memory_format=torch.channels_last)
        pyramid = self.cnn(x)
        if not upsample:
            with torch.no_grad():
                vit = self._vit(x.device)
                tok = vit.patch_tokens(x.to(vit.cls_token.dtype))          # (B, hw, 1024): already NHWC
                hp, wp = H // 14, W // 14
                pyramid[16] = tok.view(B, hp, wp, -1).permute(0, 3, 1, 2)  # (B,1024,hp,wp) channels-last view
        return pyramid
