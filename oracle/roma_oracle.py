"""CPU oracle for the RoMa matching hot path — TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Only `tests/`, `__graft_entry__.smoke()` and the `cpu_baseline` / `--impl reference` legs of `bench.py`
may import this module; it is the checker the CUDA kernels are compared against and the CPU baseline
that is timed beside them.  Nothing under `roma_b200/` imports it.

It restates, in plain CPU torch/numpy written from the definitions (explicit index arithmetic instead of
`F.grid_sample` / `F.interpolate` / `torch.cdist`), what the reference computes on this path.  Each
function cites the reference file:line it follows (paths relative to the reference repo root).

Pinning: the reference has no golden vectors or tests of its own.  The oracle is
pinned against outputs of the REFERENCE ITSELF, executed in the build container by
`tests/golden/make_golden.py` (which imports /root/reference) and committed as small fixtures under
`tests/golden/`; `tests/test_oracle_golden.py` checks every function below against them.
"""
from __future__ import annotations

import math
from typing import Optional, Tuple

import torch
import torch.nn.functional as F

# The explicit-index restatements below are what parity tests compare against.  For the TIMED CPU baseline
# (bench.py cpu_baseline / --impl reference) the same functions can instead issue the ATen library calls the
# reference itself issues at those call sites (F.grid_sample, torch.cdist), so that the baseline has the
# reference's CPU performance characteristics rather than those of a gather-by-index_select restatement.
# tests/test_oracle_golden.py checks that both modes agree.
USE_ATEN = False


def set_aten(flag: bool) -> None:
    global USE_ATEN
    USE_ATEN = bool(flag)


# ----------------------------------------------------------------------------------------------
# coordinate conventions
# ----------------------------------------------------------------------------------------------
def pixel_centers(n: int, device=None, dtype=torch.float32) -> torch.Tensor:
    """linspace(-1 + 1/n, 1 - 1/n, n): pixel-centre coordinates (matcher.py:113-114, 236-237, 699-700)."""
    return torch.linspace(-1 + 1 / n, 1 - 1 / n, n, device=device, dtype=dtype)


def pixel_grid(b: int, h: int, w: int, device=None) -> torch.Tensor:
    """(b,2,h,w) grid with channel 0 = x, channel 1 = y (matcher.py:111-118)."""
    ys, xs = pixel_centers(h, device), pixel_centers(w, device)
    g = torch.stack((xs[None, :].expand(h, w), ys[:, None].expand(h, w)))
    return g[None].expand(b, 2, h, w)


# ----------------------------------------------------------------------------------------------
# A3: bilinear grid_sample, zeros padding, align_corners=False
# ----------------------------------------------------------------------------------------------
def grid_sample_zeros(inp: torch.Tensor, grid: torch.Tensor) -> torch.Tensor:
    """F.grid_sample(inp, grid, mode='bilinear', padding_mode='zeros', align_corners=False) — matcher.py:109.

    inp (B,C,H,W), grid (B,h,w,2) normalised (x, y).  ATen semantics: x = ((g+1)*W - 1)/2, four integer
    neighbours, each weighted by the opposite-corner area, neighbours outside the image contribute 0.
    """
    if USE_ATEN:
        return F.grid_sample(inp.float(), grid.float(), mode="bilinear", padding_mode="zeros", align_corners=False)
    B, C, H, W = inp.shape
    _, h, w, _ = grid.shape
    src = inp.permute(0, 2, 3, 1).reshape(B * H * W, C).float()
    gx, gy = grid[..., 0].float(), grid[..., 1].float()
    ix = ((gx + 1) * W - 1) / 2
    iy = ((gy + 1) * H - 1) / 2
    x0, y0 = torch.floor(ix), torch.floor(iy)
    wx1, wy1 = ix - x0, iy - y0
    wx0, wy0 = (x0 + 1) - ix, (y0 + 1) - iy
    base = (torch.arange(B, device=inp.device) * (H * W)).view(B, 1, 1)
    out = torch.zeros(B * h * w, C, dtype=torch.float32, device=inp.device)
    for dy, wy in ((0, wy0), (1, wy1)):
        for dx, wx in ((0, wx0), (1, wx1)):
            xx, yy = x0 + dx, y0 + dy
            ok = (xx >= 0) & (xx <= W - 1) & (yy >= 0) & (yy <= H - 1)
            idx = base + yy.clamp(0, H - 1).long() * W + xx.clamp(0, W - 1).long()
            wgt = (wx * wy * ok).reshape(-1, 1)
            out += src.index_select(0, idx.reshape(-1)) * wgt
    return out.view(B, h, w, C).permute(0, 3, 1, 2)


# ----------------------------------------------------------------------------------------------
# A4: local correlation
# ----------------------------------------------------------------------------------------------
def local_correlation(feature0: torch.Tensor, feature1: torch.Tensor, local_radius: int,
                      flow: Optional[torch.Tensor] = None) -> torch.Tensor:
    """romatch/utils/local_correlation.py:4-48 (padding zeros, bilinear).

    corr[b,k,i,j] = sum_c f0[b,c,i,j]/sqrt(C) * bilinear(f1[b,c], flow[b,:,i,j] + offset_k),
    k = ky*(2r+1)+kx, offset_k = (linspace(-2r/w, 2r/w, 2r+1)[kx], linspace(-2r/h, 2r/h, 2r+1)[ky]).
    Every tap is sampled independently from its own normalised coordinate, exactly as the reference
    does (the CUDA kernel's shared-fraction shortcut is NOT used here).
    """
    r = int(local_radius)
    B, C, h, w = feature0.shape
    K1 = 2 * r + 1
    if flow is None:
        flow = pixel_grid(B, h, w, feature0.device)
    coords = flow.permute(0, 2, 3, 1).float()
    oy = torch.linspace(-2 * r / h, 2 * r / h, K1, device=feature0.device)
    ox = torch.linspace(-2 * r / w, 2 * r / w, K1, device=feature0.device)
    f0 = feature0.float() / (C ** 0.5)
    f1 = feature1.float()
    if USE_ATEN:
        # the reference's formulation (local_correlation.py:30-47): per batch element ONE grid_sample over all
        # (2r+1)^2 shifted copies of the coordinate field, then a channel reduction against f0
        off = torch.stack(torch.meshgrid(oy, ox, indexing="ij")[::-1], dim=-1).reshape(1, 1, K1 * K1, 2)    # (x, y)
        out = []
        for b in range(B):
            g = (coords[b][:, :, None, :] + off).reshape(1, h, w * K1 * K1, 2)
            win = F.grid_sample(f1[b:b + 1], g, mode="bilinear", padding_mode="zeros", align_corners=False)
            out.append((f0[b, :, :, :, None] * win.reshape(C, h, w, K1 * K1)).sum(dim=0).permute(2, 0, 1))
        return torch.stack(out)
    corr = torch.empty(B, K1 * K1, h, w, dtype=torch.float32, device=feature0.device)
    for ky in range(K1):
        for kx in range(K1):
            off = torch.stack((ox[kx], oy[ky])).view(1, 1, 1, 2)
            corr[:, ky * K1 + kx] = (f0 * grid_sample_zeros(f1, coords + off)).sum(dim=1)
    return corr


# ----------------------------------------------------------------------------------------------
# A5: ConvRefiner input assembly
# ----------------------------------------------------------------------------------------------
def refiner_inputs(x: torch.Tensor, y: torch.Tensor, flow: torch.Tensor,
                   disp_w: Optional[torch.Tensor], disp_b: Optional[torch.Tensor],
                   disp_scale: float, local_corr_radius: Optional[int]) -> torch.Tensor:
    """d = cat(x, x_hat, disp_emb(40/32*scale_factor*(flow - grid)), local_corr) — matcher.py:105-132.

    disp_scale is 40/32*scale_factor (matcher.py:120); disp_w (E,2[,1,1]) / disp_b (E) the 1x1 conv."""
    B, C, h, w = x.shape
    parts = [x.float(), grid_sample_zeros(y, flow.permute(0, 2, 3, 1))]
    if disp_w is not None:
        E = disp_w.shape[0]
        disp = disp_scale * (flow.float() - pixel_grid(B, h, w, x.device))
        emb = torch.einsum("ed,bdhw->behw", disp_w.reshape(E, 2).float(), disp) + disp_b.float().view(1, E, 1, 1)
        parts.append(emb)
    if local_corr_radius:
        parts.append(local_correlation(x, y, local_corr_radius, flow))
    return torch.cat(parts, dim=1)


# ----------------------------------------------------------------------------------------------
# A6 / A7: update and bilinear resampling
# ----------------------------------------------------------------------------------------------
def _resize_axis(n_in: int, n_out: int, device):
    """ATen upsample_bilinear2d source index, align_corners=False: src=(dst+0.5)*in/out-0.5 clamped at 0."""
    scale = torch.tensor(n_in / n_out, dtype=torch.float32, device=device)
    src = (torch.arange(n_out, dtype=torch.float32, device=device) + 0.5) * scale - 0.5
    src = src.clamp(min=0)
    i0 = src.floor().long().clamp(max=n_in - 1)
    i1 = i0 + (i0 < n_in - 1).long()
    l1 = src - i0.float()
    return i0, i1, l1


def bilinear_resize(t: torch.Tensor, size: Tuple[int, int]) -> torch.Tensor:
    """F.interpolate(t, size=size, mode='bilinear', align_corners=False) — matcher.py:349-360, 407-417, 657."""
    t = t.float()
    H, W = t.shape[-2:]
    h2, w2 = int(size[0]), int(size[1])
    y0, y1, ly = _resize_axis(H, h2, t.device)
    x0, x1, lx = _resize_axis(W, w2, t.device)
    ly, lx = ly.view(-1, 1), lx.view(1, -1)
    r0, r1 = t.index_select(-2, y0), t.index_select(-2, y1)
    top = (1 - lx) * r0.index_select(-1, x0) + lx * r0.index_select(-1, x1)
    bot = (1 - lx) * r1.index_select(-1, x0) + lx * r1.index_select(-1, x1)
    return (1 - ly) * top + ly * bot


def flow_update(flow: torch.Tensor, certainty, delta: torch.Tensor, ins: int, H: int, W: int):
    """matcher.py:397-402: flow += ins*(dx/(4W), dy/(4H)); certainty += dcert (refine_init = 4)."""
    disp = ins * torch.stack((delta[:, 0].float() / (4 * W), delta[:, 1].float() / (4 * H)), dim=1)
    flow = flow + disp
    certainty = certainty + delta[:, 2:3].float()
    return flow, certainty


# ----------------------------------------------------------------------------------------------
# A8: match epilogue
# ----------------------------------------------------------------------------------------------
def match_epilogue(flow: torch.Tensor, certainty: torch.Tensor, cert16: Optional[torch.Tensor], symmetric: bool):
    """matcher.py:656-662 and 684-730.  Returns (warp (B,H,Wout,4), certainty (B,H,Wout), logits (B,H,Wout))."""
    Bd, _, H, W = flow.shape
    logits = certainty.float()
    if cert16 is not None:
        low = bilinear_resize(cert16, (H, W))
        low = 0.5 * low * (low < 0)
        logits = logits - low
    f = flow.float().permute(0, 2, 3, 1)
    prob = logits.sigmoid()
    wrong = (f.abs() > 1).sum(dim=-1) > 0
    prob = torch.where(wrong[:, None], torch.zeros_like(prob), prob)
    f = f.clamp(-1, 1)
    B = Bd // 2 if symmetric else Bd
    grid = pixel_grid(B, H, W, flow.device).permute(0, 2, 3, 1)
    if symmetric:
        a2b, b2a = f.chunk(2)
        warp = torch.cat((torch.cat((grid, a2b), dim=-1), torch.cat((b2a, grid), dim=-1)), dim=2)
        prob = torch.cat(prob.chunk(2), dim=3)
        logits = torch.cat(logits.chunk(2), dim=3)
    else:
        warp = torch.cat((grid, f), dim=-1)
    return warp, prob[:, 0], logits[:, 0]


# ----------------------------------------------------------------------------------------------
# A1 / A2: GP coarse matcher
# ----------------------------------------------------------------------------------------------
def cos_kernel(x: torch.Tensor, y: torch.Tensor, T: float = 0.2, eps: float = 1e-6) -> torch.Tensor:
    """CosKernel.__call__ — matcher.py:154-163.  x (b,n,d), y (b,m,d) -> (b,n,m)."""
    c = torch.einsum("bnd,bmd->bnm", x, y) / (x.norm(dim=-1)[..., None] * y.norm(dim=-1)[:, None] + eps)
    return ((c - 1.0) / T).exp()


def gp_pos_basis(b: int, h: int, w: int, pos_w: torch.Tensor, pos_b: torch.Tensor) -> torch.Tensor:
    """f = cos(8*pi*pos_conv(grid)) as (b, h*w, D) — matcher.py:232-247 + :224."""
    D = pos_w.shape[0]
    g = pixel_grid(b, h, w, pos_w.device).to(pos_w.dtype)
    lin = torch.einsum("ed,bdhw->behw", pos_w.reshape(D, 2), g) + pos_b.view(1, D, 1, 1)
    return torch.cos(8 * math.pi * lin).flatten(2).transpose(1, 2)


def gp_posterior(x: torch.Tensor, y: torch.Tensor, pos_w: torch.Tensor, pos_b: torch.Tensor,
                 T: float = 0.2, sigma_noise: float = 0.1, eps: float = 1e-6,
                 dtype: torch.dtype = torch.float32) -> torch.Tensor:
    """GP.forward with no_cov=True — matcher.py:249-273: mu = K_xy (K_yy + sigma I)^-1 f -> (b,D,h1,w1).

    dtype=float32 follows the reference op for op (explicit inverse, matcher.py:259-263); dtype=float64
    gives a higher-precision comparand for error budgets."""
    b, c, h1, w1 = x.shape
    _, _, h2, w2 = y.shape
    xf = x.float().flatten(2).transpose(1, 2).to(dtype)
    yf = y.float().flatten(2).transpose(1, 2).to(dtype)
    f = gp_pos_basis(b, h2, w2, pos_w.to(dtype), pos_b.to(dtype))
    K_yy = cos_kernel(yf, yf, T, eps)
    K_xy = cos_kernel(xf, yf, T, eps)
    A = K_yy + sigma_noise * torch.eye(h2 * w2, dtype=dtype, device=x.device)[None]
    # one LAPACK getrf/getri per element: the same arithmetic as the reference's batched torch.linalg.inv
    # (matcher.py:261), which dead-locks inside oneMKL (SLASWP) on this torch CPU build for batch > 1 and n >~ 100
    K_yy_inv = torch.stack([torch.linalg.inv(a) for a in A])
    mu = K_xy.matmul(K_yy_inv.matmul(f))
    return mu.transpose(1, 2).reshape(b, -1, h1, w1).float()


# ----------------------------------------------------------------------------------------------
# A10 / A9: kde and balanced sampling
# ----------------------------------------------------------------------------------------------
def _euclid_dist(x1: torch.Tensor, x2: torch.Tensor) -> torch.Tensor:
    """What torch.cdist(p=2) computes for > 25 rows (ATen _euclidean_dist): |a|^2 + |b|^2 - 2ab by ONE matmul
    over augmented vectors, clamp_min(0), sqrt — in the dtype of the inputs (kde.py:9-11 calls it in fp16)."""
    x1n = x1.pow(2).sum(dim=-1, keepdim=True)
    x2n = x2.pow(2).sum(dim=-1, keepdim=True)
    a = torch.cat((x1.mul(-2), x1n, torch.ones_like(x1n)), dim=-1)
    b = torch.cat((x2, torch.ones_like(x2n), x2n), dim=-1)
    return a.matmul(b.mT).clamp_min(0).sqrt()


def kde(x: torch.Tensor, std: float = 0.1, half: bool = True, down: Optional[int] = None,
        chunk: int = 4096) -> torch.Tensor:
    """romatch/utils/kde.py:4-12: density_i = sum_j exp(-|x_i - x_j|^2 / (2 std^2)), in fp16 when `half`.

    Row-chunked so N=40000 does not need an N x N temporary; each chunk follows the reference op chain
    (cdist -> **2 -> neg -> / (2 std^2) -> exp -> sum) in the same dtype."""
    if half:
        x = x.half()
    ref = x if down is None else x[::down]
    if USE_ATEN:
        return (-torch.cdist(x, ref) ** 2 / (2 * std ** 2)).exp().sum(dim=-1)          # kde.py:9-12 verbatim op chain
    out = torch.empty(x.shape[0], dtype=x.dtype, device=x.device)
    for s in range(0, x.shape[0], chunk):
        d = _euclid_dist(x[s:s + chunk], ref)
        out[s:s + chunk] = (-d ** 2 / (2 * std ** 2)).exp().sum(dim=-1)
    return out


def sample(matches: torch.Tensor, certainty: torch.Tensor, num: int = 10000, sample_thresh: float = 0.05,
           sample_mode: str = "threshold_balanced", generator: Optional[torch.Generator] = None,
           kde_fn=None, return_indices: bool = False):
    """RegressionMatcher.sample — matcher.py:468-495 (and tiny.py:297-327 via kde_fn).

    `generator` makes the two torch.multinomial draws reproducible; with return_indices the flat int64
    indices into matches.reshape(-1,4) are returned as a third value."""
    if "threshold" in sample_mode:
        certainty = certainty.clone()
        certainty[certainty > sample_thresh] = 1
    matches, certainty = matches.reshape(-1, 4), certainty.reshape(-1)
    expansion = 4 if "balanced" in sample_mode else 1
    good = torch.multinomial(certainty, num_samples=min(expansion * num, len(certainty)), replacement=False,
                             generator=generator)
    good_matches, good_certainty = matches[good], certainty[good]
    if "balanced" not in sample_mode:
        return (good_matches, good_certainty, good) if return_indices else (good_matches, good_certainty)
    density = (kde_fn or kde)(good_matches, std=0.1)
    p = 1 / (density + 1)
    p[density < 10] = 1e-7
    bal = torch.multinomial(p, num_samples=min(num, len(good_certainty)), replacement=False, generator=generator)
    out = (good_matches[bal], good_certainty[bal])
    return out + (good[bal],) if return_indices else out


# ----------------------------------------------------------------------------------------------
# A11: match_keypoints
# ----------------------------------------------------------------------------------------------
def match_keypoints(x_A: torch.Tensor, x_B: torch.Tensor, warp: torch.Tensor, certainty: torch.Tensor,
                    sample_thresh: float = 0.05):
    """matcher.py:576-591 with return_inds=True: mutual nearest neighbours between warp(x_A) and x_B."""
    xa2b = grid_sample_zeros(warp[..., -2:].permute(2, 0, 1)[None], x_A[None, None])[0, :, 0].mT
    cert = grid_sample_zeros(certainty[None, None], x_A[None, None])[0, 0, 0]
    D = (xa2b[:, None, :] - x_B[None, :, :]).pow(2).sum(-1).sqrt()
    mutual = (D == D.min(dim=-1, keepdim=True).values) & (D == D.min(dim=-2, keepdim=True).values) \
        & (cert[:, None] > sample_thresh)
    return torch.nonzero(mutual, as_tuple=True)


# ----------------------------------------------------------------------------------------------
# (f-1): cls_to_flow_refine
# ----------------------------------------------------------------------------------------------
def cls_to_flow_refine(cls: torch.Tensor) -> torch.Tensor:
    """romatch/utils/utils.py:301-323: softmax over res^2 bins, mode, 5-tap probability-weighted mean -> (B,H,W,2)."""
    B, C, H, W = cls.shape
    res = round(math.sqrt(C))
    c = pixel_centers(res, cls.device)
    G = torch.stack((c[None, :].expand(res, res), c[:, None].expand(res, res)), dim=-1).reshape(C, 2)
    p = cls.float().softmax(dim=1)
    mode = p.max(dim=1).indices
    index = torch.stack((mode - 1, mode, mode + 1, mode - res, mode + res), dim=1).clamp(0, C - 1)
    nb = torch.gather(p, 1, index)[..., None]
    flow = sum(nb[:, i] * G[index[:, i]] for i in range(5))
    return flow / nb.sum(dim=1)


# ----------------------------------------------------------------------------------------------
# T1 / T2: Tiny-RoMa correlation volume and approximate soft-argmax
# ----------------------------------------------------------------------------------------------
def tiny_corr_volume(feat0: torch.Tensor, feat1: torch.Tensor) -> torch.Tensor:
    """romatch/models/tiny.py:241-254 -> (B,H1,W1,H0,W0)."""
    B, C, H0, W0 = feat0.shape
    _, _, H1, W1 = feat1.shape
    cv = torch.einsum("bci,bcj->bji", feat0.reshape(B, C, -1), feat1.reshape(B, C, -1))
    return cv.reshape(B, H1, W1, H0, W0) / math.sqrt(C)


def tiny_pos_embed(corr_volume: torch.Tensor) -> torch.Tensor:
    """romatch/models/tiny.py:178-199 (inference branch): argmax over all image-1 positions; softmax over the
    4x4-strided subset PLUS one extra logit equal to float(argmax index) (tiny.py:196); expectation -> (B,2,H0,W0)."""
    B, H1, W1, H0, W0 = corr_volume.shape
    down = 4
    dev = corr_volume.device
    xs, ys = pixel_centers(W1, dev), pixel_centers(H1, dev)
    grid = torch.stack((xs[None, :].expand(H1, W1), ys[:, None].expand(H1, W1)), dim=-1).reshape(H1 * W1, 2)
    xl = torch.linspace(-1 + down / W1, 1 - down / W1, W1 // down, device=dev)
    yl = torch.linspace(-1 + down / H1, 1 - down / H1, H1 // down, device=dev)
    grid_lr = torch.stack((xl[None, :].expand(H1 // down, W1 // down), yl[:, None].expand(H1 // down, W1 // down)),
                          dim=-1).reshape(-1, 2)
    best = corr_volume.reshape(B, H1 * W1, H0, W0).argmax(dim=1)
    sub = corr_volume[:, ::down, ::down].reshape(B, -1, H0, W0)
    P = torch.cat((sub, best[:, None].to(sub.dtype)), dim=1).softmax(dim=1)
    pos = torch.einsum("bchw,cd->bdhw", P[:, :-1], grid_lr)
    return pos + P[:, -1:] * grid[best].permute(0, 3, 1, 2)
