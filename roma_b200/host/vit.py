"""Plain-PyTorch ViT bodies that stay PyTorch (north_star): the DINOv2 ViT-L/14 feature encoder and the
5-block "embedding decoder" transformer.

Parameter names mirror the reference so the same checkpoints load unchanged:
  DINOv2       romatch/models/transformer/dinov2.py:43-358 (cls_token, pos_embed, mask_token, patch_embed.proj,
               blocks.N.{norm1, attn.qkv, attn.proj, ls1.gamma, norm2, mlp.fc1, mlp.fc2, ls2.gamma}, norm)
  Block        romatch/models/transformer/layers/block.py:36-105, attention.py:29-81, mlp.py, layer_scale.py
Attention runs through F.scaled_dot_product_attention (library flash kernel on the GPU).
"""
from __future__ import annotations

import math
from typing import Optional

import torch
import torch.nn as nn
import torch.nn.functional as F


def _radd(x: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
    """Residual add; the fp32-stream + 16-bit-branch case of the embedding decoder under autocast runs as one vectorised kernel."""
    if x.is_cuda and x.dtype == torch.float32 and y.dtype in (torch.float16, torch.bfloat16) and x.shape == y.shape and x.numel() % 4 == 0:
        from .. import ops
        return ops.residual_add(x, y)
    return x + y


class RowNorm(nn.LayerNorm):
    """nn.LayerNorm (same parameters / state_dict keys) that runs `roma_layernorm` on CUDA at inference."""
    cast_under_autocast = False      # set on the norms whose only consumer is a Linear (Block.norm1 / norm2)

    def native(self, x: torch.Tensor) -> bool:
        D = self.normalized_shape[0]
        return (x.is_cuda and not self.training and x.dtype in (torch.float16, torch.bfloat16, torch.float32)
                and D % 8 == 0 and D <= (1024 if x.dtype == torch.float32 else 2048))

    def forward(self, x: torch.Tensor, shift: Optional[torch.Tensor] = None) -> torch.Tensor:
        if self.native(x):
            from .. import ops
            if (x.dtype == torch.float32 and shift is None and self.cast_under_autocast and torch.is_autocast_enabled()
                    and torch.get_autocast_gpu_dtype() in (torch.float16, torch.bfloat16) and self.normalized_shape[0] <= 1024):
                # fp32 token stream under autocast (the embedding decoder): the consumer is a Linear, which would cast the fp32
                # result to the autocast dtype first — write that dtype directly (same bits, one pass over the tokens less)
                return ops.layer_norm(x, self.weight, self.bias, self.eps, out_dtype=torch.get_autocast_gpu_dtype())
            return ops.layer_norm(x, self.weight, self.bias, self.eps, shift=shift)
        return super().forward(x if shift is None else (x.float() + shift).to(x.dtype))


class SelfAttention(nn.Module):
    def __init__(self, dim: int, num_heads: int, qkv_bias: bool, proj_bias: bool = True):
        super().__init__()
        self.num_heads = num_heads
        self.qkv = nn.Linear(dim, dim * 3, bias=qkv_bias)
        self.proj = nn.Linear(dim, dim, bias=proj_bias)

    def attend(self, x: torch.Tensor) -> torch.Tensor:
        """Attention output BEFORE the output projection, (B, N, C)."""
        B, N, C = x.shape
        q, k, v = self.qkv(x).view(B, N, 3, self.num_heads, C // self.num_heads).permute(2, 0, 3, 1, 4)
        o = F.scaled_dot_product_attention(q, k, v)          # softmax(q k^T / sqrt(d)) v
        return o.transpose(1, 2).reshape(B, N, C)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return self.proj(self.attend(x))


def gelu_(h: torch.Tensor) -> torch.Tensor:
    """nn.GELU() (mlp.py:37) on a fresh fc1 output: 16-bit CUDA tensors go through `roma_gelu` in place (one streaming
    pass at HBM speed; correctly rounded exact GELU), everything else through ATen."""
    if (h.is_cuda and h.dtype in (torch.bfloat16, torch.float16) and h.is_contiguous() and h.numel() % 8 == 0
            and not (torch.is_grad_enabled() and h.requires_grad)):
        from .. import ops
        return ops.gelu(h, inplace=True)
    return F.gelu(h)


class FeedForward(nn.Module):
    def __init__(self, dim: int, hidden: int, bias: bool = True):
        super().__init__()
        self.fc1 = nn.Linear(dim, hidden, bias=bias)
        self.fc2 = nn.Linear(hidden, dim, bias=bias)

    def forward(self, x):
        return self.fc2(gelu_(self.fc1(x)))


class ChannelScale(nn.Module):
    def __init__(self, dim: int, init: float):
        super().__init__()
        self.gamma = nn.Parameter(init * torch.ones(dim))

    def forward(self, x):
        return x * self.gamma


class Block(nn.Module):
    """Pre-norm transformer block; `init_values` adds DINOv2's per-channel residual scale."""

    def __init__(self, dim: int, num_heads: int, mlp_ratio: float = 4.0, qkv_bias: bool = False,
                 init_values: Optional[float] = None, norm_eps: float = 1e-5):
        super().__init__()
        self.norm1 = RowNorm(dim, eps=norm_eps)
        self.norm1.cast_under_autocast = True           # consumed by attn.qkv only
        self.attn = SelfAttention(dim, num_heads, qkv_bias)
        self.ls1 = ChannelScale(dim, init_values) if init_values else nn.Identity()
        self.norm2 = RowNorm(dim, eps=norm_eps)
        self.norm2.cast_under_autocast = True           # consumed by mlp.fc1 only
        self.mlp = FeedForward(dim, int(dim * mlp_ratio))
        self.ls2 = ChannelScale(dim, init_values) if init_values else nn.Identity()
        self._fold_key, self._fold = None, None

    def _folded(self):
        """LayerScale folded into the linear layer it follows (gamma * (W o + b) = (gamma W) o + gamma b): at
        inference this removes two full-tensor multiplies per block.  Cached per parameter version."""
        ps = (self.attn.proj.weight, self.attn.proj.bias, self.ls1.gamma, self.mlp.fc2.weight, self.mlp.fc2.bias, self.ls2.gamma)
        key = tuple((p.data_ptr(), p._version, p.dtype) for p in ps)
        if self._fold_key != key:
            with torch.inference_mode(False), torch.no_grad():   # cached tensors must be normal (not inference) tensors
                dt = ps[0].dtype
                g1, g2 = self.ls1.gamma.float(), self.ls2.gamma.float()
                self._fold = ((ps[0].float() * g1[:, None]).to(dt), (ps[1].float() * g1).to(dt),
                              (ps[3].float() * g2[:, None]).to(dt), (ps[4].float() * g2).to(dt))
            self._fold_key = key
        return self._fold

    def foldable(self, x) -> bool:
        return (not self.training and x.is_cuda and isinstance(self.ls1, ChannelScale) and self.attn.proj.bias is not None
                and self.norm1.native(x))

    def forward_unbiased(self, x, shift):
        """x is the residual stream WITHOUT the bias vectors accumulated so far (`shift`, fp32 (D)): both residual
        branches are GEMMs accumulating into x in place (beta = 1: no separate add, no bias pass); the bias re-enters in
        fp32 inside the next LayerNorm (`roma_layernorm_shift`).  Returns (x, shift + b1 + b2); x is modified in place."""
        w1, b1, w2, b2 = self._folded()
        B, N, D = x.shape
        x2 = x.view(B * N, D)
        x2.addmm_(self.attn.attend(self.norm1(x, shift)).reshape(B * N, D), w1.t())
        shift = shift + b1.float()
        x2.addmm_(gelu_(self.mlp.fc1(self.norm2(x, shift))).reshape(B * N, -1), w2.t())
        return x, shift + b2.float()

    def fusable(self, x) -> bool:
        """16-bit CUDA inference with LayerScale: the residual adds can ride the LayerNorm that follows them."""
        return self.foldable(x) and x.dtype in (torch.float16, torch.bfloat16) and x.is_contiguous()

    def forward_deferred(self, x, pending, next_norm_of=None):
        """One block with its residual adds fused into the following norms (`roma_add_layernorm`: sum and LayerNorm(sum) in
        one pass, same bits as add + norm).  `pending` is the previous block's second branch, not yet added to x (None for
        the first block).  Returns (x, pending'): the stream after this block's FIRST add and this block's second branch."""
        from .. import ops
        w1, b1, w2, b2 = self._folded()
        if pending is None:
            n1 = self.norm1(x)
        else:
            x, n1 = ops.add_layer_norm(x, pending, self.norm1.weight, self.norm1.bias, self.norm1.eps, inplace=True)
        a = F.linear(self.attn.attend(n1), w1, b1)
        x, n2 = ops.add_layer_norm(x, a, self.norm2.weight, self.norm2.bias, self.norm2.eps, inplace=True)
        return x, F.linear(gelu_(self.mlp.fc1(n2)), w2, b2)

    def forward(self, x):
        if not self.training and x.is_cuda and isinstance(self.ls1, ChannelScale) and self.attn.proj.bias is not None:
            w1, b1, w2, b2 = self._folded()
            x = _radd(x, F.linear(self.attn.attend(self.norm1(x)), w1, b1))
            return _radd(x, F.linear(gelu_(self.mlp.fc1(self.norm2(x))), w2, b2))
        x = _radd(x, self.ls1(self.attn(self.norm1(x))))
        return _radd(x, self.ls2(self.mlp(self.norm2(x))))


class PatchEmbed(nn.Module):
    """dinov2 patch_embed.proj (layers/patch_embed.py): a stride-14 14x14 convolution == one GEMM over non-overlapping
    patches.  At inference on CUDA it runs as that GEMM (patches unfolded by a reshape, K = 3*14*14 = 588 padded to 592
    so that rows are 16-byte aligned): cuDNN's 3-input-channel convolution took 0.92 ms per 16 x 560 x 560 call, the GEMM
    takes the time of the 30 MB unfold copy."""

    def __init__(self, patch: int, in_chans: int, dim: int):
        super().__init__()
        self.proj = nn.Conv2d(in_chans, dim, kernel_size=patch, stride=patch)
        self._wcache = None

    def _gemm_weight(self):
        w = self.proj.weight
        key = (w.device, w._version, w.data_ptr())
        if self._wcache is None or self._wcache[0] != key:
            K = w[0].numel()
            Kp = (K + 7) // 8 * 8
            with torch.inference_mode(False), torch.no_grad():
                wp = torch.zeros(w.shape[0], Kp, device=w.device, dtype=w.dtype)
                wp[:, :K] = w.detach().reshape(w.shape[0], K)
            self._wcache = (key, wp)
        return self._wcache[1]

    def forward(self, x):
        p = self.proj.kernel_size[0]
        B, C, H, W = x.shape
        if self.training or not x.is_cuda or H % p or W % p:
            return self.proj(x).flatten(2).transpose(1, 2)
        wp = self._gemm_weight()
        K, hp, wq = C * p * p, H // p, W // p
        cols = x.new_zeros((B, hp * wq, wp.shape[1])) if wp.shape[1] != K else x.new_empty((B, hp * wq, K))
        cols[:, :, :K].view(B, hp, wq, C, p, p).copy_(x.view(B, C, hp, p, wq, p).permute(0, 2, 4, 1, 3, 5))
        return F.linear(cols, wp, self.proj.bias)


class DinoViT(nn.Module):
    """DINOv2 ViT (defaults = ViT-L/14 as built by encoders.py:86-95: img 518, patch 14, LayerScale 1.0, MLP FFN)."""

    def __init__(self, img_size: int = 518, patch_size: int = 14, embed_dim: int = 1024, depth: int = 24,
                 num_heads: int = 16, init_values: float = 1.0, init: bool = True):
        super().__init__()
        self.patch_size = patch_size
        self.patch_embed = PatchEmbed(patch_size, 3, embed_dim)
        n = (img_size // patch_size) ** 2
        self.cls_token = nn.Parameter(torch.zeros(1, 1, embed_dim))
        self.pos_embed = nn.Parameter(torch.zeros(1, n + 1, embed_dim))
        self.blocks = nn.ModuleList(Block(embed_dim, num_heads, 4.0, qkv_bias=True, init_values=init_values,
                                          norm_eps=1e-6) for _ in range(depth))
        self.norm = RowNorm(embed_dim, eps=1e-6)
        self.mask_token = nn.Parameter(torch.zeros(1, embed_dim))
        if init:
            nn.init.normal_(self.pos_embed, std=0.02)
            nn.init.normal_(self.cls_token, std=1e-6)
            for m in self.modules():
                if isinstance(m, nn.Linear):
                    nn.init.normal_(m.weight, std=0.02)
                    if m.bias is not None:
                        nn.init.zeros_(m.bias)
        self.requires_grad_(False)
        self._pos_cache = {}
        self.fused_residual_norm = True  # residual adds ride the LayerNorm that follows them (Block.forward_deferred; same bits)
        self.unbiased_residual = False   # opt-in: residual branches as in-place accumulating GEMMs (Block.forward_unbiased);
                                         # measured neutral on the power-capped step (95.9 vs 95.2 / 96.3 ms), so off by default

    def _pos(self, hp: int, wp: int, dtype) -> torch.Tensor:
        """Position table resampled to an hp x wp patch grid (dinov2.py:166-191: bicubic, +0.1 scale nudge)."""
        N = self.pos_embed.shape[1] - 1
        side = int(math.sqrt(N))
        if hp * wp == N and hp == wp:
            return self.pos_embed.to(dtype)
        key = (hp, wp, dtype, self.pos_embed.device, self.pos_embed._version)
        hit = self._pos_cache.get(key)
        if hit is not None:
            return hit
        pe = self.pos_embed.float()
        grid = pe[:, 1:].reshape(1, side, side, -1).permute(0, 3, 1, 2)
        grid = F.interpolate(grid, scale_factor=((hp + 0.1) / side, (wp + 0.1) / side), mode="bicubic")
        assert grid.shape[-2:] == (hp, wp)
        out = torch.cat((pe[:, :1], grid.permute(0, 2, 3, 1).reshape(1, hp * wp, -1)), dim=1).to(dtype)
        self._pos_cache = {key: out}
        return out

    def patch_tokens(self, x: torch.Tensor) -> torch.Tensor:
        """forward_features(x)['x_norm_patchtokens'] (dinov2.py:225-241): (B,3,H,W) -> (B, H/14*W/14, dim)."""
        B, _, H, W = x.shape
        t = self.patch_embed(x)
        t = torch.cat((self.cls_token.expand(B, -1, -1).to(t.dtype), t), dim=1)
        t = t + self._pos(H // self.patch_size, W // self.patch_size, t.dtype)
        if self.unbiased_residual and t.is_contiguous() and all(blk.foldable(t) for blk in self.blocks):
            shift = torch.zeros(t.shape[-1], device=t.device, dtype=torch.float32)
            for blk in self.blocks:
                t, shift = blk.forward_unbiased(t, shift)
            return self.norm(t, shift)[:, 1:]
        if self.fused_residual_norm and not torch.is_grad_enabled() and all(blk.fusable(t) for blk in self.blocks) and self.norm.native(t):
            from .. import ops
            pending = None                                                  # t is a fresh tensor: updated in place from here on
            for blk in self.blocks:
                t, pending = blk.forward_deferred(t, pending)
            return ops.add_layer_norm(t, pending, self.norm.weight, self.norm.bias, self.norm.eps, inplace=True)[1][:, 1:]
        for blk in self.blocks:
            t = blk(t)
        return self.norm(t)[:, 1:]
