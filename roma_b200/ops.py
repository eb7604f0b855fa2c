"""Tensor-level wrappers over the C ABI (include/roma_b200.h), with the reference's call signatures.

PyTorch is plumbing here: it owns device memory and the stream; every computation below is one of the
hand-written sm_100a kernels.  Nothing in this module falls back to ATen or the CPU — tensors that are
not on a CUDA device raise.
"""
from __future__ import annotations

import contextlib
import ctypes as C
import math
import threading
from typing import Optional, Tuple

import torch

from . import _lib
from ._lib import Strides4

_DT = {torch.float32: _lib.ROMA_F32, torch.float16: _lib.ROMA_F16, torch.bfloat16: _lib.ROMA_BF16}


class KernelTimer:
    """Optional per-launch CUDA-event timing of the C-ABI calls (used by bench.py for the roofline block).

    with KernelTimer() as t: ...   records (name, dims, start_event, end_event) for every kernel call made on the
    current stream inside the block; t.summary() synchronises and returns {name: [(dims, ms), ...]}."""
    active = None

    def __init__(self):
        self.records = []

    def __enter__(self):
        KernelTimer.active = self
        return self

    def __exit__(self, *exc):
        KernelTimer.active = None

    def summary(self):
        torch.cuda.synchronize()
        out = {}
        for name, dims, s, e in self.records:
            out.setdefault(name, []).append((dims, s.elapsed_time(e)))
        return out


_tls = threading.local()


def _run(name: str, dims: dict, call):
    """Invoke one C-ABI entry point (optionally bracketed by events on the current stream) and check its code.

    The launch happens with the TENSORS' device current (recorded by `_cuda` for this thread): the library launches on
    the calling thread's current device, and `_stream(dev)` / the pointers belong to the tensors' device — without the
    guard a call on cuda:1 from a process whose current device is 0 would launch on GPU 0 against GPU-1 memory."""
    dev = getattr(_tls, "dev", None)
    guard = torch.cuda.device(dev) if dev is not None and dev.index is not None and dev.index != torch.cuda.current_device() \
        else contextlib.nullcontext()
    with guard:
        t = KernelTimer.active
        if t is None:
            rc = call()
        else:
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            rc = call()
            e.record()
            t.records.append((name, dims, s, e))
    _lib.check(rc, name)


def _cuda(*ts: Optional[torch.Tensor]) -> torch.device:
    dev = None
    for t in ts:
        if t is None:
            continue
        if not t.is_cuda:
            raise RuntimeError("roma_b200 kernels run on sm_100a only; got a tensor on "
                               f"{t.device} (there is no CPU fallback)")
        if dev is None:
            dev = t.device
        elif t.device != dev:
            raise RuntimeError(f"tensors on different devices: {dev} vs {t.device}")
    _tls.dev = dev
    return dev


def _ptr(t: Optional[torch.Tensor]) -> C.c_void_p:
    return C.c_void_p(0 if t is None else t.data_ptr())


def _stream(dev: torch.device) -> C.c_void_p:
    return C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)


def _s4(t: Optional[torch.Tensor]) -> Strides4:
    """Element strides of a (b, c, y, x)-indexed tensor."""
    if t is None:
        return Strides4(0, 0, 0, 0)
    assert t.dim() == 4
    return Strides4(*t.stride())


def _nhwc(t: torch.Tensor) -> torch.Tensor:
    """(B,C,H,W) tensor -> same tensor laid out channels-last and dense (no copy if it already is)."""
    if t.dim() != 4:
        raise ValueError(f"expected a (B,C,H,W) tensor, got {tuple(t.shape)}")
    return t.contiguous(memory_format=torch.channels_last)


def _f32(t: torch.Tensor) -> torch.Tensor:
    return t if t.dtype == torch.float32 else t.float()


# --------------------------------------------------------------------------------------------------
# K2
# --------------------------------------------------------------------------------------------------
def refiner_inputs(x: torch.Tensor, y: Optional[torch.Tensor], flow: torch.Tensor,
                   disp_w: Optional[torch.Tensor] = None, disp_b: Optional[torch.Tensor] = None,
                   disp_scale: float = 1.0, local_corr_radius: Optional[int] = None, pad_to: int = 1,
                   swap_halves: bool = False) -> torch.Tensor:
    """ConvRefiner input d = cat(x, x_hat, disp_emb, local_corr) in ONE kernel (matcher.py:105-132).

    x, y: (B,C,h,w) features of image A / B (any of fp32/fp16/bf16, same dtype); flow: (B,2,h,w) fp32 view.
    Returns d as a (B, 2C+E+K, h, w) tensor in channels-last memory format, dtype of x.  With pad_to > 1 the
    channel count is rounded up to a multiple of pad_to and the extra channels are zero: rows become 16-byte
    aligned (pad_to=8 for 16-bit features), which is what the vectorised kernel path and the tensor-core GEMMs of
    the refiner body want; the caller pads its weights accordingly.
    swap_halves=True (symmetric matching, matcher.py:459-463): y is x with the two halves of the batch exchanged;
    no swapped copy is made — the kernel runs once per half on views (y may be None).
    """
    dev = _cuda(x, y, flow, disp_w, disp_b)
    if swap_halves:
        y = x
    if x.dtype != y.dtype or x.dtype not in _DT:
        raise ValueError(f"unsupported feature dtypes {x.dtype}/{y.dtype}")
    B, Cc, h, w = x.shape
    _, _, h1, w1 = y.shape
    x, y, flow = _nhwc(x), _nhwc(y), _f32(flow)
    r = int(local_corr_radius or 0)
    E = 0 if disp_w is None else disp_w.shape[0]
    K = (2 * r + 1) ** 2 if r > 0 else 0
    if E:
        disp_w = disp_w.reshape(E, 2).float().contiguous()
        disp_b = disp_b.reshape(E).float().contiguous()
    dch = 2 * Cc + E + K
    if pad_to > 1:
        dch = (dch + pad_to - 1) // pad_to * pad_to
    d = torch.empty((B, dch, h, w), device=dev, dtype=x.dtype, memory_format=torch.channels_last)
    lib = _lib.load()
    if swap_halves:
        if B % 2:
            raise ValueError("swap_halves needs an even batch")
        parts = [(x[:B // 2], x[B // 2:], flow[:B // 2], d[:B // 2]), (x[B // 2:], x[:B // 2], flow[B // 2:], d[B // 2:])]
    else:
        parts = [(x, y, flow, d)]
    for xp, yp, fp_, dp in parts:
        nws = lib.roma_refiner_gather_workspace_bytes(xp.shape[0], h, w, r)      # tile map of the shared-patch kernel (r = 2, 3)
        ws = torch.empty(nws, device=dev, dtype=torch.uint8) if nws else None
        _run("roma_refiner_gather", dict(B=xp.shape[0], C=Cc, h=h, w=w, r=r, E=E, esize=x.element_size()),
             lambda: lib.roma_refiner_gather_ws(
                 _ptr(xp), _ptr(yp), _ptr(fp_), _s4(fp_), _ptr(disp_w), _ptr(disp_b), float(disp_scale),
                 _ptr(dp), dch, 0, Cc, (2 * Cc if E else -1), (2 * Cc + E if r > 0 else -1),
                 xp.shape[0], Cc, h, w, h1, w1, E, r, _DT[x.dtype], _ptr(ws), nws, _stream(dev)))
    return d


def local_correlation(feature0: torch.Tensor, feature1: torch.Tensor, local_radius: int,
                      padding_mode: str = "zeros", flow: Optional[torch.Tensor] = None,
                      sample_mode: str = "bilinear") -> torch.Tensor:
    """Drop-in for romatch.utils.local_correlation.local_correlation (local_correlation.py:4-48)."""
    if padding_mode != "zeros" or sample_mode != "bilinear":
        raise NotImplementedError("roma_b200.local_correlation implements padding_mode='zeros', sample_mode='bilinear'")
    dev = _cuda(feature0, feature1, flow)
    B, Cc, h, w = feature0.shape
    r = int(local_radius)
    if tuple(feature1.shape[2:]) != (h, w):
        raise NotImplementedError(f"roma_b200.local_correlation needs feature maps of one size, got {tuple(feature0.shape)} and "
                                  f"{tuple(feature1.shape)} (the reference's window step is then not one pixel of feature1)")
    if flow is None:   # aligned features: sample around the pixel's own centre (local_correlation.py:16-27)
        ys = torch.linspace(-1 + 1 / h, 1 - 1 / h, h, device=dev)
        xs = torch.linspace(-1 + 1 / w, 1 - 1 / w, w, device=dev)
        flow = torch.stack((xs[None, :].expand(h, w), ys[:, None].expand(h, w)))[None].expand(B, 2, h, w)
    f0, f1, flow = _nhwc(feature0), _nhwc(feature1), _f32(flow)
    K = (2 * r + 1) ** 2
    out = torch.empty((B, K, h, w), device=dev, dtype=feature0.dtype, memory_format=torch.channels_last)
    lib = _lib.load()
    _run("roma_refiner_gather", dict(B=B, C=Cc, h=h, w=w, r=r, E=0, esize=f0.element_size(), only="corr"),
         lambda: lib.roma_refiner_gather(
             _ptr(f0), _ptr(f1), _ptr(flow), _s4(flow), _ptr(None), _ptr(None), 0.0,
             _ptr(out), K, -1, -1, -1, 0,
             B, Cc, h, w, feature1.shape[2], feature1.shape[3], 0, r, _DT[feature0.dtype], _stream(dev)))
    return out


def grid_sample(input: torch.Tensor, grid: torch.Tensor) -> torch.Tensor:
    """F.grid_sample(input, grid, mode='bilinear', padding_mode='zeros', align_corners=False) (matcher.py:109)."""
    dev = _cuda(input, grid)
    B, Cc, h1, w1 = input.shape
    _, h, w, two = grid.shape
    assert two == 2
    f1 = _nhwc(input)
    flow = _f32(grid).permute(0, 3, 1, 2)   # view: (B,2,h,w) with arbitrary strides, consumed in place
    out = torch.empty((B, Cc, h, w), device=dev, dtype=input.dtype, memory_format=torch.channels_last)
    lib = _lib.load()
    _run("roma_refiner_gather", dict(B=B, C=Cc, h=h, w=w, r=0, E=0, esize=f1.element_size(), only="xhat"),
         lambda: lib.roma_refiner_gather(
             _ptr(None), _ptr(f1), _ptr(flow), _s4(flow), _ptr(None), _ptr(None), 0.0,
             _ptr(out), Cc, -1, 0, -1, -1,
             B, Cc, h, w, h1, w1, 0, 0, _DT[input.dtype], _stream(dev)))
    return out


# --------------------------------------------------------------------------------------------------
# K3a
# --------------------------------------------------------------------------------------------------
def flow_update_resample(flow: torch.Tensor, certainty: Optional[torch.Tensor], delta: Optional[torch.Tensor],
                         ins: int, full_hw: Tuple[int, int], next_hw: Optional[Tuple[int, int]],
                         want_current: bool = True):
    """flow += ins*(delta_xy/(4*W,4*H)); certainty += delta_c; then bilinear resample to next_hw.

    matcher.py:397-402 and :349-360 / :407-417.  flow (B,2,h,w), certainty (B,1,h,w) or None, delta (B,3,h,w) or None.
    Returns (flow_cur, cert_cur, flow_next, cert_next) as (B,2,·,·)/(B,1,·,·) views over interleaved storage;
    entries not requested are None.
    """
    dev = _cuda(flow, certainty, delta)
    B, _, h, w = flow.shape
    flow = _f32(flow)
    certainty = None if certainty is None else _f32(certainty)
    delta = None if delta is None else _f32(delta)
    H, W = full_hw
    fo = co = fn = cn = None
    if want_current:
        fo = torch.empty((B, h, w, 2), device=dev, dtype=torch.float32)
        co = torch.empty((B, h, w), device=dev, dtype=torch.float32)
    h2 = w2 = 0
    if next_hw is not None:
        h2, w2 = int(next_hw[0]), int(next_hw[1])
        fn = torch.empty((B, h2, w2, 2), device=dev, dtype=torch.float32)
        cn = torch.empty((B, h2, w2), device=dev, dtype=torch.float32)
    lib = _lib.load()
    _run("roma_flow_update_resample", dict(B=B, h=h, w=w, h2=h2, w2=w2, cur=bool(want_current), delta=delta is not None),
         lambda: lib.roma_flow_update_resample(
             _ptr(flow), _s4(flow), _ptr(certainty), _s4(certainty), _ptr(delta), _s4(delta),
             float(ins), float(4 * W), float(4 * H),
             _ptr(fo), _ptr(co), _ptr(fn), _ptr(cn), B, h, w, h2, w2, _stream(dev)))
    v = lambda f: None if f is None else f.permute(0, 3, 1, 2)
    u = lambda c: None if c is None else c[:, None]
    return v(fo), u(co), v(fn), u(cn)


def match_epilogue(flow: torch.Tensor, certainty: torch.Tensor, cert16: Optional[torch.Tensor],
                   symmetric: bool, return_logits: bool = False):
    """matcher.py:656-662 + 684-718 in one pass.  Returns (warp (B,H,Wout,4), certainty (B,H,Wout)[, logits])."""
    dev = _cuda(flow, certainty, cert16)
    Bd, _, H, W = flow.shape
    if symmetric and Bd % 2:
        raise ValueError("symmetric epilogue needs an even decoder batch")
    B = Bd // 2 if symmetric else Bd
    Wout = 2 * W if symmetric else W
    flow, certainty = _f32(flow), _f32(certainty)
    cert16 = None if cert16 is None else _f32(cert16)
    warp = torch.empty((B, H, Wout, 4), device=dev, dtype=torch.float32)
    cert = torch.empty((B, H, Wout), device=dev, dtype=torch.float32)
    logits = torch.empty((B, H, Wout), device=dev, dtype=torch.float32) if return_logits else None
    h16, w16 = (cert16.shape[2], cert16.shape[3]) if cert16 is not None else (0, 0)
    lib = _lib.load()
    _run("roma_match_epilogue", dict(B=B, H=H, W=W, Wout=Wout, logits=bool(return_logits)),
         lambda: lib.roma_match_epilogue(
             _ptr(flow), _s4(flow), _ptr(certainty), _s4(certainty), _ptr(cert16), _s4(cert16), h16, w16,
             _ptr(warp), _ptr(cert), _ptr(logits), B, H, W, int(bool(symmetric)), _stream(dev)))
    return (warp, cert, logits) if return_logits else (warp, cert)


# --------------------------------------------------------------------------------------------------
# K1
# --------------------------------------------------------------------------------------------------
_gp_ws = {}


def gp_posterior(x: torch.Tensor, y: torch.Tensor, pos_w: torch.Tensor, pos_b: torch.Tensor,
                 T: float = 0.2, sigma_noise: float = 0.1, eps: float = 1e-6) -> torch.Tensor:
    """GP.forward with CosKernel, fourier basis, no_cov=True (matcher.py:249-273).

    x (B,C,h1,w1), y (B,C,h2,w2) features (fp32/fp16/bf16); returns mu (B,D,h1,w1) fp32 (channels-last)."""
    dev = _cuda(x, y, pos_w, pos_b)
    if x.dtype != y.dtype or x.dtype not in _DT:
        raise ValueError(f"unsupported feature dtypes {x.dtype}/{y.dtype}")
    B, Cc, h1, w1 = x.shape
    _, _, h2, w2 = y.shape
    D = pos_w.shape[0]
    x, y = _nhwc(x), _nhwc(y)
    pos_w = pos_w.reshape(D, 2).float().contiguous()
    pos_b = pos_b.reshape(D).float().contiguous()
    lib = _lib.load()
    need = lib.roma_gp_workspace_bytes(B, h1 * w1, h2 * w2, Cc, D)
    key = (dev, torch.cuda.current_stream(dev).cuda_stream)
    ws = _gp_ws.get(key)
    if ws is None or ws.numel() < need:
        ws = torch.empty(need, device=dev, dtype=torch.uint8)
        _gp_ws[key] = ws
    mu = torch.empty((B, D, h1, w1), device=dev, dtype=torch.float32, memory_format=torch.channels_last)
    _run("roma_gp_posterior", dict(B=B, Nx=h1 * w1, Ny=h2 * w2, C=Cc, D=D),
         lambda: lib.roma_gp_posterior(_ptr(x), _ptr(y), _ptr(pos_w), _ptr(pos_b), _ptr(mu),
                                       B, h1 * w1, h2 * w2, Cc, D, h2, w2, float(T), float(sigma_noise), float(eps),
                                       _DT[x.dtype], _ptr(ws), ws.numel(), _stream(dev)))
    return mu


def tc_gemm_nt(a: torch.Tensor, b: torch.Tensor, split: Optional[int] = None) -> torch.Tensor:
    """Batched a @ b^T on the tcgen05 path of K1: a (batch,M,K), b (batch,N,K) -> (batch,M,N) fp32."""
    dev = _cuda(a, b)
    if a.dtype != b.dtype or a.dtype not in _DT:
        raise ValueError(f"unsupported dtypes {a.dtype}/{b.dtype}")
    a, b = a.contiguous(), b.contiguous()
    batch, M, K = a.shape
    N = b.shape[1]
    split = split or (3 if a.dtype == torch.float32 else 1)
    out = torch.empty((batch, M, N), device=dev, dtype=torch.float32)
    lib = _lib.load()
    _run("roma_tc_gemm_nt", dict(batch=batch, M=M, N=N, K=K, split=split),
         lambda: lib.roma_tc_gemm_nt(_ptr(a), _ptr(b), _ptr(out), batch, M, N, K, _DT[a.dtype], split, _stream(dev)))
    return out


# --------------------------------------------------------------------------------------------------
# K3b: kde + sampling
# --------------------------------------------------------------------------------------------------
def kde(x: torch.Tensor, std: float = 0.1, half: bool = True, down: Optional[int] = None,
        return_p: bool = False):
    """Drop-in for romatch.utils.kde.kde (kde.py:4-12): streaming, no N x N temporary.

    Returns the density as fp16 when `half` (like the reference) else fp32; with return_p also the sampling
    weights p = 1/(density+1), p[density<10] = 1e-7 of matcher.py:490-491 in the same dtype."""
    dev = _cuda(x)
    if x.dim() != 2 or x.shape[1] != 4:
        raise ValueError(f"kde expects (N,4) points, got {tuple(x.shape)}")
    N = x.shape[0]
    xf = x.float().contiguous()
    lib = _lib.load()
    dens = torch.empty(N, device=dev, dtype=torch.float32)
    p = torch.empty(N, device=dev, dtype=torch.float32) if return_p else None
    ws = torch.empty(lib.roma_kde_workspace_bytes(N), device=dev, dtype=torch.uint8)
    _run("roma_kde", dict(N=N, down=int(down or 1), half=bool(half)),
         lambda: lib.roma_kde(_ptr(xf), N, int(down or 1), float(std), int(bool(half)), _ptr(dens), _ptr(p),
                              _ptr(ws), ws.numel(), _stream(dev)))
    if half:   # values are already fp16-representable: the cast is exact
        dens = dens.half()
        p = None if p is None else p.half()
    return (dens, p) if return_p else dens


_TOPK_DTYPES = {torch.float32: 0, torch.float16: 1, torch.bfloat16: 2}
# ATen's topk takes the multi-block radix select + stable radix sort of the survivors (whose tie order roma_topk_desc
# reproduces) for one slice of >= 20000 elements and more than 4096 results; smaller problems go through its single-block /
# bitonic variants, whose tie order is not claimed here: they stay on torch.topk.
_TOPK_MIN_N, _TOPK_MIN_K = 20000, 4097


def topk_indices(q: torch.Tensor, k: int) -> torch.Tensor:
    """Indices (int64, (k,)) of the k largest entries of the 1-D tensor q, ordered by (value descending, index ascending)
    — `torch.topk(q, k).indices` on CUDA for the sizes `sample` uses, ties included (tests/test_topk_gpu.py)."""
    dev = _cuda(q)
    if q.dim() != 1 or q.dtype not in _TOPK_DTYPES:
        raise ValueError(f"topk_indices: expected a 1-D fp32/fp16/bf16 tensor, got {tuple(q.shape)} {q.dtype}")
    q = q.contiguous()
    n, k = q.numel(), int(k)
    lib = _lib.load()
    out = torch.empty(k, device=dev, dtype=torch.int64)
    with torch.cuda.device(dev):
        ws = torch.empty(lib.roma_topk_workspace_bytes(n, k), device=dev, dtype=torch.uint8)
    _run("roma_topk_desc", dict(n=n, k=k),
         lambda: lib.roma_topk_desc(_ptr(q), _TOPK_DTYPES[q.dtype], n, k, _ptr(out), None, _ptr(ws), ws.numel(), _stream(dev)))
    return out


def multinomial_without_replacement(p: torch.Tensor, num_samples: int, generator: Optional[torch.Generator] = None) -> torch.Tensor:
    """torch.multinomial(p, num_samples, replacement=False) for a 1-D weight vector, restated around the own top-k:
    ATen (native/Distributions.cpp, multinomial_out) checks the weights with two device-side asserts, draws
    q ~ Exp(1) of p's shape and dtype from the generator, divides p / q in place and returns topk(p / q, num_samples).
    The same ops in the same order on the same generator (so the Philox stream advances as in the reference), with
    `topk_indices` as the last step."""
    n = p.numel()
    if p.dim() != 1 or not (_TOPK_MIN_K <= num_samples <= n and n >= _TOPK_MIN_N) or p.dtype not in _TOPK_DTYPES:
        return torch.multinomial(p, num_samples=num_samples, replacement=False, generator=generator)
    _cuda(p)
    torch._assert_async((p.max() < float("inf")) & (p.min() >= 0),
                        "probability tensor contains either `inf`, `nan` or element < 0")
    torch._assert_async(~(p.sum() == 0), "invalid multinomial distribution (sum of probabilities <= 0)")
    q = torch.empty_like(p).exponential_(1, generator=generator)
    torch.div(p, q, out=q)
    return topk_indices(q, num_samples)


def sample(matches: torch.Tensor, certainty: torch.Tensor, num: int = 10000, sample_thresh: float = 0.05,
           sample_mode: str = "threshold_balanced", kde_half: bool = True, kde_down: Optional[int] = None,
           generator: Optional[torch.Generator] = None, return_indices: bool = False):
    """RegressionMatcher.sample (matcher.py:468-495) / TinyRoMa.sample (tiny.py:297-327).

    The two draws consume the device generator's Philox stream exactly as torch.multinomial does in the reference
    (`multinomial_without_replacement`: ATen's exponential draw and division, own top-k with ATen's tie order) — that
    is what makes the sampled indices reproducible against it; thresholding is a torch elementwise op, the density
    and the weights p come from the kde kernel."""
    _cuda(matches, certainty)
    if "threshold" in sample_mode:
        certainty = torch.where(certainty > sample_thresh, torch.ones_like(certainty), certainty)
    matches, certainty = matches.reshape(-1, 4), certainty.reshape(-1)
    expansion = 4 if "balanced" in sample_mode else 1
    good = multinomial_without_replacement(certainty, min(expansion * num, len(certainty)), generator)
    good_matches, good_certainty = matches[good], certainty[good]
    if "balanced" not in sample_mode:
        return (good_matches, good_certainty, good) if return_indices else (good_matches, good_certainty)
    _, p = kde(good_matches, std=0.1, half=kde_half, down=kde_down, return_p=True)
    bal = multinomial_without_replacement(p, min(num, len(good_certainty)), generator)
    out = (good_matches[bal], good_certainty[bal])
    return out + (good[bal],) if return_indices else out


# --------------------------------------------------------------------------------------------------
# K3c: match_keypoints
# --------------------------------------------------------------------------------------------------
def match_keypoints(x_A: torch.Tensor, x_B: torch.Tensor, warp: torch.Tensor, certainty: torch.Tensor,
                    sample_thresh: float = 0.05):
    """matcher.py:576-580: int64 (inds_A, inds_B) of mutual nearest neighbours with certainty > thresh."""
    dev = _cuda(x_A, x_B, warp, certainty)
    Na, Nb = x_A.shape[0], x_B.shape[0]
    H, Wp, four = warp.shape
    assert four == 4 and certainty.shape == (H, Wp)
    xa, xb = x_A.float().contiguous(), x_B.float().contiguous()
    warp, certainty = warp.float().contiguous(), certainty.float().contiguous()
    lib = _lib.load()
    cap = max(Na, Nb)
    ia = torch.empty(cap, device=dev, dtype=torch.int64)
    ib = torch.empty(cap, device=dev, dtype=torch.int64)
    count = torch.zeros(1, device=dev, dtype=torch.int32)
    ws = torch.empty(lib.roma_match_keypoints_workspace_bytes(Na, Nb), device=dev, dtype=torch.uint8)
    _run("roma_match_keypoints", dict(Na=Na, Nb=Nb),
         lambda: lib.roma_match_keypoints(_ptr(xa), Na, _ptr(xb), Nb, _ptr(warp), _ptr(certainty), H, Wp,
                                          float(sample_thresh), _ptr(ia), _ptr(ib), cap, _ptr(count),
                                          _ptr(ws), ws.numel(), _stream(dev)))
    n = int(count.item())   # the reference's torch.nonzero has the same inherent sync (matcher.py:580)
    if n > cap:
        raise RuntimeError(f"match_keypoints: {n} tied pairs exceed capacity {cap}")
    return ia[:n], ib[:n]


# --------------------------------------------------------------------------------------------------
# (f-2) refiner body: depthwise 5x5 + folded BN + ReLU
# --------------------------------------------------------------------------------------------------
def dwconv5x5_bias_act(x: torch.Tensor, weight_taps: torch.Tensor, bias: torch.Tensor, relu: bool = True) -> torch.Tensor:
    """Depthwise 5x5 conv (pad 2) + bias (+ ReLU) on a channels-last (B,C,h,w) tensor (matcher.py:77-103 at inference,
    BatchNorm folded by the caller).  weight_taps: fp32 (25, C) tap-major; bias: fp32 (C).  C must be even."""
    dev = _cuda(x, weight_taps, bias)
    if x.dtype not in _DT:
        raise ValueError(f"unsupported dtype {x.dtype}")
    B, Cc, h, w = x.shape
    x = _nhwc(x)
    out = torch.empty_like(x, memory_format=torch.channels_last)
    lib = _lib.load()
    _run("roma_dwconv5x5_bias_act", dict(B=B, C=Cc, h=h, w=w, esize=x.element_size()),
         lambda: lib.roma_dwconv5x5_bias_act(_ptr(x), _ptr(weight_taps), _ptr(bias), _ptr(out), B, h, w, Cc, int(bool(relu)),
                                             _DT[x.dtype], _stream(dev)))
    return out


def dwconv5x5_pointwise(x: torch.Tensor, dw_weight_taps: torch.Tensor, dw_bias: torch.Tensor, pw_weight: torch.Tensor,
                        pw_bias: torch.Tensor, relu: bool = True) -> torch.Tensor:
    """One ConvRefiner block at the finest scale's padded width (C = 24, fp16 / bf16): depthwise 5x5 + bias + ReLU and the
    1x1 convolution that follows it (matcher.py:77-103), fused — the depthwise result stays in shared memory.
    x (B,24,h,w) channels-last; dw_weight_taps fp32 (25,24); pw_weight fp32 (24,24) [out][in]."""
    dev = _cuda(x, dw_weight_taps, dw_bias, pw_weight, pw_bias)
    if x.dtype not in (torch.float16, torch.bfloat16):
        raise ValueError(f"unsupported dtype {x.dtype}")
    B, Cc, h, w = x.shape
    x = _nhwc(x)
    pw_weight = pw_weight.reshape(Cc, Cc).float().contiguous()
    pw_bias = pw_bias.float().contiguous()
    out = torch.empty_like(x, memory_format=torch.channels_last)
    lib = _lib.load()
    _run("roma_dwconv5x5_pointwise", dict(B=B, C=Cc, h=h, w=w, esize=x.element_size()),
         lambda: lib.roma_dwconv5x5_pointwise(_ptr(x), _ptr(dw_weight_taps), _ptr(dw_bias), _ptr(pw_weight), _ptr(pw_bias), _ptr(out),
                                              B, h, w, Cc, int(bool(relu)), _DT[x.dtype], _stream(dev)))
    return out


def pointwise_narrow(x: torch.Tensor, weight: torch.Tensor, bias: torch.Tensor) -> torch.Tensor:
    """1x1 convolution to OUT <= 16 channels in the dtype of x (roma_models.py:131-142, proj["1"] with its BatchNorm
    folded).  x (B,C,h,w) channels-last; weight fp32 (OUT,C); bias fp32 (OUT) -> (B,OUT,h,w) channels-last."""
    dev = _cuda(x, weight, bias)
    if x.dtype not in _DT:
        raise ValueError(f"unsupported dtype {x.dtype}")
    B, Cc, h, w = x.shape
    OUT = weight.shape[0]
    x = _nhwc(x)
    weight = weight.reshape(OUT, Cc).float().contiguous()
    bias = bias.float().contiguous()
    out = torch.empty((B, h, w, OUT), device=dev, dtype=x.dtype)
    lib = _lib.load()
    _run("roma_pointwise_narrow", dict(B=B, C=Cc, OUT=OUT, h=h, w=w, esize=x.element_size()),
         lambda: lib.roma_pointwise_narrow(_ptr(x), _ptr(weight), _ptr(bias), _ptr(out), B * h * w, Cc, OUT,
                                           _DT[x.dtype], _stream(dev)))
    return out.permute(0, 3, 1, 2)


def layer_norm(x: torch.Tensor, weight: torch.Tensor, bias: torch.Tensor, eps: float,
               shift: Optional[torch.Tensor] = None, out_dtype: Optional[torch.dtype] = None) -> torch.Tensor:
    """F.layer_norm(x [+ shift], (D,), weight, bias, eps) over the last dimension (block.py:88-101, dinov2.py:236).
    shift: optional fp32 (D) vector added in fp32 in front of the statistics (`roma_layernorm_shift`)."""
    dev = _cuda(x, weight, bias, shift)
    if x.dtype not in _DT:
        raise ValueError(f"unsupported dtype {x.dtype}")
    D = x.shape[-1]
    x = x.contiguous()
    weight, bias = weight.to(x.dtype).contiguous(), bias.to(x.dtype).contiguous()
    rows = x.numel() // D
    lib = _lib.load()
    if out_dtype is not None and out_dtype != x.dtype:
        # fp32 rows, 16-bit output: autocast's fp32 layer_norm + cast in front of the consuming Linear, in one pass
        if x.dtype != torch.float32 or out_dtype not in (torch.float16, torch.bfloat16) or shift is not None:
            raise ValueError("layer_norm(out_dtype=...) is for fp32 rows written as fp16 / bf16")
        out = torch.empty(x.shape, device=dev, dtype=out_dtype)
        _run("roma_layernorm", dict(rows=rows, D=D, esize=4, out_esize=2),
             lambda: lib.roma_layernorm_cast(_ptr(x), _ptr(weight), _ptr(bias), _ptr(out), rows, D, float(eps), _DT[out_dtype], _stream(dev)))
        return out
    out = torch.empty_like(x)
    if shift is not None:
        shift = shift.reshape(D).float().contiguous()
        _run("roma_layernorm", dict(rows=rows, D=D, esize=x.element_size()),
             lambda: lib.roma_layernorm_shift(_ptr(x), _ptr(shift), _ptr(weight), _ptr(bias), _ptr(out), rows, D, float(eps),
                                              _DT[x.dtype], _stream(dev)))
        return out
    _run("roma_layernorm", dict(rows=rows, D=D, esize=x.element_size()),
         lambda: lib.roma_layernorm(_ptr(x), _ptr(weight), _ptr(bias), _ptr(out), rows, D, float(eps), _DT[x.dtype], _stream(dev)))
    return out


def add_layer_norm(x: torch.Tensor, y: torch.Tensor, weight: torch.Tensor, bias: torch.Tensor, eps: float,
                   inplace: bool = False):
    """(x + y, F.layer_norm(x + y)) in one pass for x, y of one dtype and shape (block.py:88-101: a residual add and the norm
    that follows it).  inplace=True writes the sum over x.  Same bits as the two ATen / roma ops run one after the other."""
    dev = _cuda(x, y, weight, bias)
    if x.dtype not in _DT or y.dtype != x.dtype or x.shape != y.shape:
        raise ValueError("add_layer_norm: x and y must share dtype and shape")
    D = x.shape[-1]
    if inplace and not x.is_contiguous():
        raise ValueError("add_layer_norm(inplace=True) needs a contiguous x")
    x, y = x.contiguous(), y.contiguous()
    weight, bias = weight.to(x.dtype).contiguous(), bias.to(x.dtype).contiguous()
    rows = x.numel() // D
    total = x if inplace else torch.empty_like(x)
    out = torch.empty_like(x)
    lib = _lib.load()
    _run("roma_add_layernorm", dict(rows=rows, D=D, esize=x.element_size()),
         lambda: lib.roma_add_layernorm(_ptr(x), _ptr(y), _ptr(weight), _ptr(bias), _ptr(total), _ptr(out), rows, D, float(eps),
                                        _DT[x.dtype], _stream(dev)))
    return total, out


def residual_add(x: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
    """x + y for an fp32 stream x and a 16-bit branch y of the same shape -> fp32 (block.py:88-101 under autocast)."""
    dev = _cuda(x, y)
    if x.dtype != torch.float32 or y.dtype not in (torch.float16, torch.bfloat16) or x.shape != y.shape or x.numel() % 4:
        raise ValueError("residual_add: fp32 + fp16/bf16 tensors of one shape, numel % 4 == 0")
    x, y = x.contiguous(), y.contiguous()
    out = torch.empty_like(x)
    lib = _lib.load()
    _run("roma_residual_add", dict(n=x.numel(), esize=4),
         lambda: lib.roma_residual_add(_ptr(x), _ptr(y), _ptr(out), x.numel(), _DT[y.dtype], _stream(dev)))
    return out


def gelu(x: torch.Tensor, inplace: bool = False) -> torch.Tensor:
    """nn.GELU() (exact, erf form; mlp.py:37) as one streaming kernel; `inplace` overwrites x (a fresh GEMM output)."""
    dev = _cuda(x)
    if x.dtype not in _DT:
        raise ValueError(f"unsupported dtype {x.dtype}")
    x = x if x.is_contiguous() else x.contiguous()
    out = x if inplace else torch.empty_like(x)
    lib = _lib.load()
    _run("roma_gelu", dict(n=x.numel(), esize=x.element_size()),
         lambda: lib.roma_gelu(_ptr(x), _ptr(out), x.numel(), _DT[x.dtype], _stream(dev)))
    return out


def maxpool2x2(x: torch.Tensor) -> torch.Tensor:
    """nn.MaxPool2d(2, 2) on a channels-last (B,C,h,w) tensor (encoders.py:57-66); returns channels-last."""
    dev = _cuda(x)
    if x.dtype not in _DT:
        raise ValueError(f"unsupported dtype {x.dtype}")
    B, Cc, h, w = x.shape
    x = _nhwc(x)
    out = torch.empty((B, Cc, h // 2, w // 2), device=dev, dtype=x.dtype, memory_format=torch.channels_last)
    lib = _lib.load()
    _run("roma_maxpool2x2_nhwc", dict(B=B, C=Cc, h=h, w=w, esize=x.element_size()),
         lambda: lib.roma_maxpool2x2_nhwc(_ptr(x), _ptr(out), B, h, w, Cc, _DT[x.dtype], _stream(dev)))
    return out


def refiner_out_conv(x: torch.Tensor, weight: torch.Tensor, bias: torch.Tensor) -> torch.Tensor:
    """ConvRefiner.out_conv in fp32 from the channels-last activation (matcher.py:141), no fp32 copy of x.

    x (B,C,h,w) channels-last, any of fp32/fp16/bf16; weight fp32 (OUT<=3, C); bias fp32 (OUT).
    Returns a (B,OUT,h,w) fp32 view over (B,h,w,4) interleaved storage."""
    dev = _cuda(x, weight, bias)
    if x.dtype not in _DT:
        raise ValueError(f"unsupported dtype {x.dtype}")
    B, Cc, h, w = x.shape
    OUT = weight.shape[0]
    x = _nhwc(x)
    weight = weight.reshape(OUT, Cc).float().contiguous()
    bias = bias.float().contiguous()
    out = torch.empty((B, h, w, 4), device=dev, dtype=torch.float32)
    lib = _lib.load()
    _run("roma_refiner_out_conv", dict(B=B, C=Cc, h=h, w=w, esize=x.element_size()),
         lambda: lib.roma_refiner_out_conv(_ptr(x), _ptr(weight), _ptr(bias), _ptr(out), B * h * w, Cc, OUT,
                                           _DT[x.dtype], _stream(dev)))
    return out.permute(0, 3, 1, 2)[:, :OUT]


# --------------------------------------------------------------------------------------------------
# (f-1) cls_to_flow_refine, KT tiny
# --------------------------------------------------------------------------------------------------
def cls_to_flow_refine(cls: torch.Tensor) -> torch.Tensor:
    """romatch/utils/utils.py:301-323: (B,C=res^2,H,W) logits -> (B,H,W,2) flow.  Strided input consumed in place."""
    dev = _cuda(cls)
    if cls.dtype not in _DT:
        raise ValueError(f"unsupported dtype {cls.dtype}")
    B, Cc, H, W = cls.shape
    out = torch.empty((B, H, W, 2), device=dev, dtype=torch.float32)
    lib = _lib.load()
    _run("roma_cls_to_flow_refine", dict(B=B, C=Cc, H=H, W=W, esize=cls.element_size()),
         lambda: lib.roma_cls_to_flow_refine(_ptr(cls), _s4(cls), _DT[cls.dtype], _ptr(out), B, Cc, H, W, _stream(dev)))
    return out


_MEAN3 = (C.c_float * 3)(0.485, 0.456, 0.406)
_STD3 = (C.c_float * 3)(0.229, 0.224, 0.225)


def prepare_images(images_u8: torch.Tensor, size_hw: Tuple[int, int], mean=None, std=None) -> torch.Tensor:
    """uint8 (B,3,H,W) -> ImageNet-normalised fp32 (B,3,h,w): (u8/255 - mean)/std and, where the size differs, the
    antialiased bicubic resize of F.interpolate(..., antialias=True) (utils.py:165-261 in tensor form), one kernel."""
    dev = _cuda(images_u8)
    if images_u8.dtype != torch.uint8 or images_u8.dim() != 4 or images_u8.shape[1] != 3:
        raise ValueError("prepare_images expects a uint8 (B,3,H,W) tensor")
    x = images_u8.contiguous()
    B, _, H, W = x.shape
    h, w = int(size_hw[0]), int(size_hw[1])
    m = _MEAN3 if mean is None else (C.c_float * 3)(*[float(v) for v in mean])
    s = _STD3 if std is None else (C.c_float * 3)(*[float(v) for v in std])
    out = torch.empty((B, 3, h, w), device=dev, dtype=torch.float32)
    lib = _lib.load()
    _run("roma_prepare_images", dict(B=B, H=H, W=W, h=h, w=w),
         lambda: lib.roma_prepare_images(_ptr(x), _ptr(out), B, H, W, h, w, C.cast(m, C.c_void_p), C.cast(s, C.c_void_p),
                                         _stream(dev)))
    return out


def prepare_images_pil(images_u8: torch.Tensor, size_hw: Tuple[int, int], mean=None, std=None, channels_last: bool = False,
                       normalise: bool = True):
    """PIL `Image.resize((w, h), Image.BICUBIC)` of 8-bit RGB images on the device, bit-exact (utils.py:223-232), fused with
    ToTensor + Normalize (utils.py:234-261).

    images_u8: uint8 (B,3,H,W), or (B,H,W,3) / (H,W,3) as decoded by PIL with `channels_last=True`; any strides.
    Returns fp32 (B,3,h,w) normalised, or — `normalise=False` — the resized uint8 (B,3,h,w) itself."""
    dev = _cuda(images_u8)
    x = images_u8
    if x.dtype != torch.uint8:
        raise ValueError("prepare_images_pil expects uint8 pixels")
    if channels_last:
        if x.dim() == 3:
            x = x[None]
        x = x.permute(0, 3, 1, 2)
    if x.dim() != 4 or x.shape[1] != 3:
        raise ValueError("prepare_images_pil expects 3-channel images: (B,3,H,W), or (B,H,W,3) with channels_last=True")
    B, _, H, W = x.shape
    h, w = int(size_hw[0]), int(size_hw[1])
    m = _MEAN3 if mean is None else (C.c_float * 3)(*[float(v) for v in mean])
    s = _STD3 if std is None else (C.c_float * 3)(*[float(v) for v in std])
    out = torch.empty((B, 3, h, w), device=dev, dtype=torch.float32 if normalise else torch.uint8)
    sb, sc, sy, sx = x.stride()
    lib = _lib.load()
    _run("roma_prepare_images_pil", dict(B=B, H=H, W=W, h=h, w=w),
         lambda: lib.roma_prepare_images_pil(_ptr(x), _ptr(out) if normalise else None, None if normalise else _ptr(out),
                                             B, H, W, h, w, sb, sc, sy, sx, C.cast(m, C.c_void_p), C.cast(s, C.c_void_p),
                                             _stream(dev)))
    return out


_tiny_ws = {}


def tiny_corr_posembed(feat0: torch.Tensor, feat1: torch.Tensor, tensor_cores: Optional[bool] = None) -> torch.Tensor:
    """TinyRoMa.pos_embed(TinyRoMa.corr_volume(feat0, feat1)) (tiny.py:241-254 + 178-199), fused.

    feat0 (B,C,H0,W0), feat1 (B,C,H1,W1) -> (B,2,H0,W0); the (B,H1,W1,H0,W0) volume is never formed.
    tensor_cores: None = the tcgen05 path whenever the shapes allow (C % 4 == 0), False = the fp32 SIMT kernel."""
    dev = _cuda(feat0, feat1)
    B, Cc, H0, W0 = feat0.shape
    _, _, H1, W1 = feat1.shape
    f0 = _nhwc(feat0.float())
    f1 = _nhwc(feat1.float())
    out = torch.empty((B, 2, H0, W0), device=dev, dtype=torch.float32)
    lib = _lib.load()
    if tensor_cores is None:
        tensor_cores = Cc % 4 == 0
    if tensor_cores:
        need = lib.roma_tiny_corr_posembed_workspace_bytes(B, H0, W0, H1, W1)
        key = (dev, torch.cuda.current_stream(dev).cuda_stream)
        ws = _tiny_ws.get(key)
        if ws is None or ws.numel() < need:
            ws = torch.empty(need, device=dev, dtype=torch.uint8)
            _tiny_ws[key] = ws
        _run("roma_tiny_corr_posembed_tc", dict(B=B, C=Cc, n0=H0 * W0, n1=H1 * W1),
             lambda: lib.roma_tiny_corr_posembed_tc(_ptr(f0), _ptr(f1), _ptr(out), B, Cc, H0, W0, H1, W1, _ptr(ws), ws.numel(),
                                                    _stream(dev)))
        return out
    _run("roma_tiny_corr_posembed", dict(B=B, C=Cc, n0=H0 * W0, n1=H1 * W1),
         lambda: lib.roma_tiny_corr_posembed(_ptr(f0), _ptr(f1), _ptr(out), B, Cc, H0, W0, H1, W1, _stream(dev)))
    return out
