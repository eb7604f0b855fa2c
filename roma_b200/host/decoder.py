"""Host-side mirror of the reference decoder (romatch/models/matcher.py:17-422) with the hot ops replaced by
the sm_100a kernels in `roma_b200.ops`.  The conv / transformer bodies stay PyTorch (north_star); module and
parameter names equal the reference's so `load_state_dict` takes the same checkpoints:

  decoder.embedding_decoder.{blocks.N.*, to_out.*}     transformer/__init__.py:10-46
  decoder.gps.16.pos_conv.*                            matcher.py:165-273
  decoder.proj.{16,8,4,2,1}.{0,1}.*                    model_zoo/roma_models.py:131-142
  decoder.conv_refiner.{s}.{block1,hidden_blocks.N,out_conv,disp_emb}.*   matcher.py:17-103

Inference only: the training-mode keys of matcher.py:386-396 are not produced (`match` forces eval, :623).
There is no CPU path: the kernels raise on non-CUDA tensors.
"""
from __future__ import annotations

from typing import Dict, Optional, Sequence

import torch
import torch.nn as nn
import torch.nn.functional as F

from .. import ops
from .encoder import _autocast
from .vit import Block


class ConvRefiner(nn.Module):
    """matcher.py:17-143.  forward(x, y, flow, scale_factor, logits) -> (delta_flow (B,2,h,w), delta_certainty (B,1,h,w)).

    Parameters / buffers are the reference's (block1, hidden_blocks.N, out_conv, disp_emb).  At inference the conv
    stack runs from a *prepared* copy of them: eval-mode BatchNorm folded into the depthwise conv, channels padded
    to a multiple of 8 with zeros (rows of the channels-last activations become 16-byte aligned, so the gather
    kernel vectorises and the 1x1 convs hit aligned tensor-core GEMMs), depthwise conv + bias + ReLU as one launch
    of the (f-2) kernel and each 1x1 conv as one library GEMM with the bias in its epilogue."""

    def __init__(self, in_dim: int, hidden_dim: int, out_dim: int = 3, kernel_size: int = 5, hidden_blocks: int = 8,
                 displacement_emb_dim: Optional[int] = None, local_corr_radius: Optional[int] = None,
                 bn_momentum: float = 0.01, amp_dtype: torch.dtype = torch.float16):
        super().__init__()
        self.block1 = self._block(in_dim, hidden_dim, kernel_size, bn_momentum)
        self.hidden_blocks = nn.Sequential(*[self._block(hidden_dim, hidden_dim, kernel_size, bn_momentum)
                                             for _ in range(hidden_blocks)])
        self.out_conv = nn.Conv2d(hidden_dim, out_dim, 1, 1, 0)
        self.has_displacement_emb = bool(displacement_emb_dim)
        if self.has_displacement_emb:
            self.disp_emb = nn.Conv2d(2, displacement_emb_dim, 1, 1, 0)
        self.local_corr_radius = local_corr_radius
        self.amp_dtype = amp_dtype
        self.in_dim, self.hidden_dim, self.kernel_size = in_dim, hidden_dim, kernel_size
        self._prep = None
        self._prep_key = None
        self.fuse_narrow_block = True     # Cp = 24: depthwise + 1x1 convolution of a block in one kernel

    @staticmethod
    def _block(cin: int, cout: int, k: int, momentum: float) -> nn.Sequential:
        assert cout % cin == 0, "depthwise block needs out_dim divisible by in_dim"
        return nn.Sequential(nn.Conv2d(cin, cout, k, 1, k // 2, groups=cin), nn.BatchNorm2d(cout, momentum=momentum),
                             nn.ReLU(inplace=True), nn.Conv2d(cout, cout, 1, 1, 0))

    # ---- prepared (folded / padded) weights -------------------------------------------------------------------
    def _prepare(self, device, dtype):
        key = (device, dtype, tuple(t._version for t in self.state_dict(keep_vars=True).values()),
               tuple(t.data_ptr() for t in self.parameters()))
        if self._prep_key == key:
            return self._prep
        assert self.in_dim == self.hidden_dim and self.kernel_size == 5, "prepared path: depthwise multiplier 1, 5x5"
        C = self.hidden_dim
        Cp = (C + 7) // 8 * 8
        blocks = []
        with torch.inference_mode(False), torch.no_grad():   # cached tensors must be normal (not inference) tensors
            for blk in [self.block1, *self.hidden_blocks]:
                conv1, bn, _, conv2 = blk
                scale = bn.weight.float() / torch.sqrt(bn.running_var.float() + bn.eps)
                shift = bn.bias.float() - bn.running_mean.float() * scale
                k2 = self.kernel_size ** 2
                w_dw = torch.zeros(k2, Cp, device=device)                       # tap-major fp32 for the dwconv kernel
                b_dw = torch.zeros(Cp, device=device)
                w_dw[:, :C] = (conv1.weight.float() * scale.view(-1, 1, 1, 1)).view(C, k2).t()
                b_dw[:C] = (conv1.bias.float() if conv1.bias is not None else 0) * scale + shift
                w_pw = torch.zeros(Cp, Cp, device=device)
                b_pw = torch.zeros(Cp, device=device)
                w_pw[:C, :C] = conv2.weight.float().view(C, C)
                b_pw[:C] = conv2.bias.float()
                if Cp <= 32:      # N = K = 24: streaming kernel with fp32 weights instead of a cuBLAS GEMM
                    blocks.append((w_dw.contiguous(), b_dw, w_pw.contiguous(), b_pw))
                else:
                    blocks.append((w_dw.contiguous(), b_dw, w_pw.to(dtype).contiguous(), b_pw.to(dtype)))
            w_out = torch.zeros(self.out_conv.out_channels, Cp, device=device)
            w_out[:, :C] = self.out_conv.weight.float().view(-1, C)
            b_out = self.out_conv.bias.float().clone()
        self._prep = (Cp, blocks, w_out, b_out)
        self._prep_key = key
        return self._prep

    def _body(self, d: torch.Tensor) -> torch.Tensor:
        """d (B, Cp, h, w) channels-last -> (B, 3, h, w) fp32 out_conv output."""
        Cp, blocks, w_out, b_out = self._prepare(d.device, d.dtype)
        B, _, h, w = d.shape
        fused = self.fuse_narrow_block and Cp == 24 and d.dtype in (torch.float16, torch.bfloat16)
        for w_dw, b_dw, w_pw, b_pw in blocks:
            if fused:
                d = ops.dwconv5x5_pointwise(d, w_dw, b_dw, w_pw, b_pw, relu=True)     # whole block in one pass
                continue
            d = ops.dwconv5x5_bias_act(d, w_dw, b_dw, relu=True)                      # (f-2): dw conv + folded BN + ReLU
            if Cp <= 32:
                d = ops.pointwise_narrow(d, w_pw, b_pw)                               # conv2, N = K = 24
            else:
                d = F.linear(d.permute(0, 2, 3, 1), w_pw, b_pw).permute(0, 3, 1, 2)   # NHWC GEMM, bias in the epilogue
        return ops.refiner_out_conv(d, w_out, b_out)                                  # matcher.py:141 (fp32 math)

    def delta(self, x, y, flow, scale_factor: float = 1, swap_halves: bool = False) -> torch.Tensor:
        """The (B,3,h,w) out_conv tensor (dx, dy, dcertainty) consumed by the fused update kernel."""
        # K2: x_hat = grid_sample(y, flow), disp_emb(40/32*scale_factor*(flow-grid)), local_correlation and the
        # concat of matcher.py:109-132 in one launch, written channels-last in the feature dtype.
        dw = self.disp_emb.weight if self.has_displacement_emb else None
        db = self.disp_emb.bias if self.has_displacement_emb else None
        d = ops.refiner_inputs(x, y, flow, dw, db, 40 / 32 * scale_factor, self.local_corr_radius, pad_to=8,
                               swap_halves=swap_halves)
        return self._body(d)

    def forward(self, x, y, flow, scale_factor: float = 1, logits=None):
        d = self.delta(x, y, flow, scale_factor)
        return d[:, :-1], d[:, -1:]


class CosKernel(nn.Module):
    """Holder of the temperature (matcher.py:145-163); the kernel itself is fused into ops.gp_posterior."""

    def __init__(self, T: float):
        super().__init__()
        self.T = T


class GP(nn.Module):
    """GP coarse matcher with CosKernel, fourier basis, no covariance output (matcher.py:165-273)."""

    def __init__(self, T: float = 0.2, gp_dim: int = 512, sigma_noise: float = 0.1):
        super().__init__()
        self.K = CosKernel(T)
        self.sigma_noise = sigma_noise
        self.pos_conv = nn.Conv2d(2, gp_dim, 1, 1)
        self.dim = gp_dim

    def forward(self, x, y, **kwargs):
        return ops.gp_posterior(x, y, self.pos_conv.weight, self.pos_conv.bias, T=self.K.T, sigma_noise=self.sigma_noise)


class TransformerDecoder(nn.Module):
    """transformer/__init__.py:10-46: tokens = cat(gp_posterior, features) -> blocks -> Linear -> (cls, certainty)."""

    def __init__(self, blocks: nn.Sequential, hidden_dim: int, out_dim: int, amp_dtype: torch.dtype = torch.float16):
        super().__init__()
        self.blocks = blocks
        self.to_out = nn.Linear(hidden_dim, out_dim)
        self.hidden_dim, self.out_dim = hidden_dim, out_dim
        self.is_classifier = True
        self.amp_dtype = amp_dtype
        self._pad_key, self._pad = None, None

    def scales(self):
        return [16]

    def forward(self, gp_posterior, features, old_stuff=None, new_scale=None):
        B, _, H, W = gp_posterior.shape
        with _autocast(gp_posterior.device, self.amp_dtype):
            # both inputs are channels-last, so (B,C,H,W)->(B,HW,C) is a view
            tok = torch.cat((gp_posterior.permute(0, 2, 3, 1).reshape(B, H * W, -1),
                             features.permute(0, 2, 3, 1).reshape(B, H * W, -1).to(gp_posterior.dtype)), dim=-1)
            hid = self.blocks(tok)
            if hid.is_cuda and not self.training and self.out_dim % 8:
                # out_dim = 64^2 + 1 = 4097 is odd: cuBLAS falls back to an unaligned legacy kernel (~10x slower).
                # Pad the output features to a multiple of 8 once; the extra columns are sliced off as a view.
                wgt, bias, npad = self._padded_to_out(hid.dtype)
                out = F.linear(hid, wgt, bias)
            else:
                out, npad = self.to_out(hid), self.out_dim
        out = out.view(B, H, W, npad).permute(0, 3, 1, 2)[:, :self.out_dim]   # (B, out_dim, H, W) view, strides kept
        return out[:, :-1], out[:, -1:], None

    def _padded_to_out(self, dtype):
        w, b = self.to_out.weight, self.to_out.bias
        key = (w.data_ptr(), w._version, b._version, dtype)
        if self._pad_key != key:
            npad = (self.out_dim + 7) // 8 * 8
            with torch.inference_mode(False), torch.no_grad():   # cached tensors must be normal (not inference) tensors
                wp = torch.zeros(npad, w.shape[1], device=w.device, dtype=dtype)
                bp = torch.zeros(npad, device=w.device, dtype=dtype)
                wp[:self.out_dim] = w.to(dtype)
                bp[:self.out_dim] = b.to(dtype)
            self._pad, self._pad_key = (wp, bp, npad), key
        return self._pad


class Decoder(nn.Module):
    """Scale loop 16 -> 8 -> 4 -> 2 -> 1 (matcher.py:275-422)."""

    def __init__(self, embedding_decoder: TransformerDecoder, gps: nn.ModuleDict, proj: nn.ModuleDict,
                 conv_refiner: nn.ModuleDict, scales: Sequence[str] = ("16", "8", "4", "2", "1"),
                 amp_dtype: torch.dtype = torch.float16):
        super().__init__()
        self.embedding_decoder = embedding_decoder
        self.gps = gps
        self.proj = proj
        self.conv_refiner = conv_refiner
        self.scales = list(scales)
        self.refine_init = 4
        self.amp_dtype = amp_dtype
        self._proj_prep = {}

    def _project(self, scale: str, f: torch.Tensor) -> torch.Tensor:
        """proj[scale] = Conv2d(1x1) + BatchNorm2d (roma_models.py:131-142).  At inference the BatchNorm is an affine
        map: it is folded into the convolution once and the result runs as ONE channels-last GEMM with the bias in
        its epilogue (instead of cudnn conv + bias add + batch_norm)."""
        conv, bn = self.proj[scale]
        if self.training:
            return self.proj[scale](f)
        dtype = self.amp_dtype if f.is_cuda else torch.float32
        key = (f.device, dtype, conv.weight._version, bn.weight._version, bn.running_mean._version, bn.running_var._version,
               conv.weight.data_ptr())
        prep = self._proj_prep.get(scale)
        if prep is None or prep[0] != key:
            with torch.inference_mode(False), torch.no_grad():   # cached tensors must be normal (not inference) tensors
                g = bn.weight.float() / torch.sqrt(bn.running_var.float() + bn.eps)
                w32 = (conv.weight.float().view(conv.out_channels, conv.in_channels) * g[:, None]).contiguous()
                b0 = conv.bias.float() if conv.bias is not None else 0.0
                b32 = ((b0 - bn.running_mean.float()) * g + bn.bias.float()).contiguous()
            prep = (key, w32.to(dtype), b32.to(dtype), w32, b32)
            self._proj_prep[scale] = prep
        _, wgt, bias, w32, b32 = prep
        if f.is_cuda and conv.out_channels <= 16 and conv.in_channels % 8 == 0:
            return ops.pointwise_narrow(f.to(dtype), w32, b32)           # N = 9: a streaming kernel, not a GEMM
        return F.linear(f.permute(0, 2, 3, 1).to(dtype), wgt, bias).permute(0, 3, 1, 2)

    def forward(self, f1: Dict[int, torch.Tensor], f2: Optional[Dict[int, torch.Tensor]], gt_warp=None, gt_prob=None,
                upsample: bool = False, flow: Optional[torch.Tensor] = None,
                certainty: Optional[torch.Tensor] = None, scale_factor: float = 1, swapped: bool = False):
        """`swapped=True` (symmetric matching, matcher.py:455-466): the support pyramid is f1 with the two halves of
        the batch exchanged; f2 may then be None and every projection is evaluated once instead of twice."""
        coarse_scales = self.embedding_decoder.scales()
        all_scales = self.scales if not upsample else ["8", "4", "2", "1"]
        sizes = {s: f1[s].shape[-2:] for s in f1}
        h, w = sizes[1]
        corresps = {}
        if upsample:
            # matcher.py:349-360: previous pass' finest flow / certainty resampled to the coarsest size of this pass
            _, _, flow, certainty = ops.flow_update_resample(flow, certainty, None, 1, (h, w),
                                                             tuple(sizes[int(all_scales[0])]), want_current=False)
        else:
            flow, certainty = None, None
        for new_scale in all_scales:
            ins = int(new_scale)
            f1_s = f1[ins]
            f2_s = None if swapped else f2[ins]
            if new_scale in self.proj:
                f1_s = self._project(new_scale, f1_s)
                f2_s = None if swapped else self._project(new_scale, f2_s)
            if ins in coarse_scales:
                if swapped:      # the GP takes whole tensors: one small swapped copy at the coarsest scale only
                    a, b = f1_s.chunk(2)
                    f2_s = torch.cat((b, a), dim=0)
                    if f1_s.is_contiguous(memory_format=torch.channels_last):
                        f2_s = f2_s.contiguous(memory_format=torch.channels_last)
                gp_posterior = self.gps[new_scale](f1_s, f2_s)                                  # K1
                cls, certainty, _ = self.embedding_decoder(gp_posterior, f1_s, None, new_scale)
                flow = ops.cls_to_flow_refine(cls).permute(0, 3, 1, 2)                          # (f-1)
            if new_scale in self.conv_refiner:
                delta = self.conv_refiner[new_scale].delta(f1_s, f2_s, flow, scale_factor=scale_factor, swap_halves=swapped)
            else:
                delta = None
            nxt = tuple(sizes[ins // 2]) if new_scale != "1" else None
            # K3a: flow += ins*(delta/(4w,4h)); certainty += dcert; bilinear resample to the next scale
            flow_s, cert_s, flow, certainty = ops.flow_update_resample(flow, certainty, delta, ins, (h, w), nxt)
            corresps[ins] = {"certainty": cert_s, "flow": flow_s}
        return corresps


def build_decoder(amp_dtype: torch.dtype = torch.float16) -> Decoder:
    """Hyper-parameters of model_zoo/roma_models.py:23-152 (roma_outdoor / roma_indoor)."""
    gp_dim = feat_dim = 512
    dec_dim = gp_dim + feat_dim
    emb = TransformerDecoder(nn.Sequential(*[Block(dec_dim, 8) for _ in range(5)]), dec_dim, 64 ** 2 + 1, amp_dtype)
    spec = {   # scale: (feature C, disp emb dim, radius)
        "16": (512, 128, 7), "8": (512, 64, 3), "4": (256, 32, 2), "2": (64, 16, None), "1": (9, 6, None)}
    refiners = nn.ModuleDict()
    for s, (c, e, r) in spec.items():
        dim = 2 * c + e + ((2 * r + 1) ** 2 if r else 0)
        refiners[s] = ConvRefiner(dim, dim, 3, 5, 8, e, r, 0.01, amp_dtype)
    gps = nn.ModuleDict({"16": GP(0.2, gp_dim)})
    proj = nn.ModuleDict({s: nn.Sequential(nn.Conv2d(ci, co, 1, 1), nn.BatchNorm2d(co))
                          for s, (ci, co) in {"16": (1024, 512), "8": (512, 512), "4": (256, 256),
                                              "2": (128, 64), "1": (64, 9)}.items()})
    return Decoder(emb, gps, proj, refiners, amp_dtype=amp_dtype)
