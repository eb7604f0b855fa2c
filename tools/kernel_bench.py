"""Per-kernel timing at the BASELINE.json config shapes (CUDA events, L2 flushed between iterations).

    python tools/kernel_bench.py [--batch 16] [--dtype bf16] [--iters 20] [--only gather,gp,...]

Prints one JSON line per measurement: algorithmic bytes / FLOPs (SURVEY.md §8d formulas), time, achieved
GB/s or TFLOP/s and the fraction of the measured peak in MEASURED_PEAKS.json.  Not a bench.py substitute:
this is the tool used to pick what to optimise and to name kernels for ncu.
"""
from __future__ import annotations

import argparse
import json
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

from roma_b200 import ops  # noqa: E402

PEAKS = {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0}
try:
    PEAKS.update(json.loads((ROOT / "MEASURED_PEAKS.json").read_text()))
except Exception:
    pass

DT = {"bf16": torch.bfloat16, "fp16": torch.float16, "fp32": torch.float32}


def timeit(fn, iters, flush):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    total = 0.0
    for _ in range(iters):
        flush.add_(1.0)                      # > L2 (126 MB): evicts the previous iteration's lines
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); fn(); e.record()
        torch.cuda.synchronize()
        total += s.elapsed_time(e)
    return total / iters


def flows(B, h, w, dev, mode, g):
    ys = torch.linspace(-1 + 1 / h, 1 - 1 / h, h); xs = torch.linspace(-1 + 1 / w, 1 - 1 / w, w)
    grid = torch.stack((xs[None, :].expand(h, w), ys[:, None].expand(h, w)))[None].expand(B, 2, h, w)
    if mode == "coherent":
        low = torch.randn(B, 2, max(h // 8, 2), max(w // 8, 2), generator=g) * 0.05
        f = grid + torch.nn.functional.interpolate(low, size=(h, w), mode="bilinear", align_corners=False)
    else:
        f = torch.rand(B, 2, h, w, generator=g) * 2.2 - 1.1
    return f.contiguous().to(dev)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=16, help="decoder batch (2 per pair)")
    ap.add_argument("--dtype", default="bf16")
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--only", default="")
    ap.add_argument("--local-corr-path", type=int, default=0, help="0 shared-patch tiles + per-pixel tensor-core kernel, 1 SIMT, 2 per-pixel tensor-core kernel only")
    ap.add_argument("--kde-mode", type=int, default=0, help="roma_kde_set_direct: 0 exact (default), 3 tensor-core every pair twice, 4 every pair once")
    ap.add_argument("--scales", default="", help="comma list of refiner visits for --only gather (s16,s8,...,s1u)")
    args = ap.parse_args()
    dev = torch.device("cuda:0")
    from roma_b200 import _lib
    _lib.load().roma_local_corr_set_path(args.local_corr_path)
    dt = DT[args.dtype]
    es = torch.tensor([], dtype=dt).element_size()
    B = args.batch
    only = set(filter(None, args.only.split(",")))
    want = lambda k: not only or k in only
    flush = torch.zeros(256 * 1024 * 1024 // 4, device=dev)
    g = torch.Generator().manual_seed(7)

    def emit(**kw):
        print(json.dumps(kw), flush=True)

    if want("gather"):
        # (name, C, E, r, h, w): the 9 refiner visits of roma_outdoor 560 -> 864
        shapes = [("s16", 512, 128, 7, 40, 40), ("s8", 512, 64, 3, 70, 70), ("s4", 256, 32, 2, 140, 140),
                  ("s2", 64, 16, 0, 280, 280), ("s1", 9, 6, 0, 560, 560), ("s8u", 512, 64, 3, 108, 108),
                  ("s4u", 256, 32, 2, 216, 216), ("s2u", 64, 16, 0, 432, 432), ("s1u", 9, 6, 0, 864, 864)]
        keep = set(filter(None, args.scales.split(",")))
        for name, C, E, r, h, w in shapes:
            if keep and name not in keep:
                continue
            x = torch.randn(B, C, h, w, device=dev, dtype=dt).contiguous(memory_format=torch.channels_last)
            y = torch.randn(B, C, h, w, device=dev, dtype=dt).contiguous(memory_format=torch.channels_last)
            dw, db = torch.randn(E, 2, device=dev), torch.randn(E, device=dev)
            K = (2 * r + 1) ** 2 if r else 0
            # f0 (if r>0... we also copy it) + f1 + flow + outputs actually written (x, x_hat, emb, corr)
            alg = B * h * w * (es * C + es * C + 8 + es * (2 * C + E + K))
            for mode in ("coherent", "incoherent"):
                fl = flows(B, h, w, dev, mode, g)
                ms = timeit(lambda: ops.refiner_inputs(x, y, fl, dw, db, 1.9, r or None, pad_to=8), args.iters, flush)
                gbs = alg / ms / 1e6
                emit(kernel="refiner_gather", scale=name, C=C, r=r, hw=[h, w], batch=B, dtype=args.dtype, flow=mode,
                     ms=round(ms, 4), alg_MB=round(alg / 1e6, 2), GBps=round(gbs, 1), frac_hbm=round(gbs / PEAKS["hbm_gbs"], 4))
            del x, y

    if want("resample"):
        for h, w, h2, w2 in [(40, 40, 70, 70), (280, 280, 560, 560), (560, 560, 108, 108), (432, 432, 864, 864)]:
            fl = flows(B, h, w, dev, "coherent", g); ce = torch.randn(B, 1, h, w, device=dev)
            de = torch.randn(B, h, w, 3, device=dev).permute(0, 3, 1, 2)
            alg = B * (h * w * (8 + 4 + 12 + 12) + h2 * w2 * 12)
            ms = timeit(lambda: ops.flow_update_resample(fl, ce, de, 2, (864, 864), (h2, w2)), args.iters, flush)
            emit(kernel="flow_update_resample", hw=[h, w], to=[h2, w2], batch=B, ms=round(ms, 4), alg_MB=round(alg / 1e6, 2),
                 GBps=round(alg / ms / 1e6, 1), frac_hbm=round(alg / ms / 1e6 / PEAKS["hbm_gbs"], 4))

    if want("epilogue"):
        H = W = 864
        fl = flows(B, H, W, dev, "coherent", g); ce = torch.randn(B, 1, H, W, device=dev); c16 = torch.randn(B, 1, 40, 40, device=dev)
        alg = B * H * W * (8 + 4 + 16 + 4)
        ms = timeit(lambda: ops.match_epilogue(fl, ce, c16, True), args.iters, flush)
        emit(kernel="match_epilogue", hw=[H, W], batch=B, ms=round(ms, 4), alg_MB=round(alg / 1e6, 2),
             GBps=round(alg / ms / 1e6, 1), frac_hbm=round(alg / ms / 1e6 / PEAKS["hbm_gbs"], 4))

    if want("gp"):
        for hw in (40, 48):
            N, C, D = hw * hw, 512, 512
            base = torch.nn.functional.interpolate(torch.randn(B, C, 5, 5, generator=g), size=(hw, hw), mode="bilinear")
            x = (base + 0.3 * torch.randn(B, C, hw, hw, generator=g)).to(dev, dt).contiguous(memory_format=torch.channels_last)
            y = (base + 0.3 * torch.randn(B, C, hw, hw, generator=g)).to(dev, dt).contiguous(memory_format=torch.channels_last)
            pw, pb = torch.randn(D, 2, device=dev), torch.randn(D, device=dev)
            flops = B * (2 * (2 * N * N * C) + N ** 3 / 3 + 2 * N * N * D + 2 * N * N * D)
            ms = timeit(lambda: ops.gp_posterior(x, y, pw, pb), max(3, args.iters // 4), flush)
            emit(kernel="gp_posterior", N=N, batch=B, dtype=args.dtype, ms=round(ms, 3), alg_GF=round(flops / 1e9, 2),
                 TFLOPs=round(flops / ms / 1e9, 2), frac_tensor=round(flops / ms / 1e9 / PEAKS["bf16_tflops"], 5))

    if want("kde"):
        from roma_b200 import _lib as _l
        _l.load().roma_kde_set_direct(args.kde_mode)
        for N in (20000, 40000):
            # rows (xA, yA, xB, yB) as sample() draws them from a symmetric warp: a third in a dense cluster, a smooth
            # correspondence field (the points lie near a 2-D manifold of [-1, 1]^4, like real candidates)
            a = torch.rand(N, 2, generator=g) * 2 - 1
            a[:N // 3] = (torch.randn(N // 3, 2, generator=g) * 0.05 + torch.tensor([0.3, -0.2])).clamp(-1, 1)
            x = torch.cat((a, (a + 0.1 * torch.sin(3 * a.flip(1)) + 0.002 * torch.randn(N, 2, generator=g)).clamp(-1, 1)), dim=1)
            x = x[torch.randperm(N, generator=g)].contiguous().to(dev)
            for half in (True, False):
                ms = timeit(lambda: ops.kde(x, half=half), max(3, args.iters // 4), flush)
                emit(kernel="kde", N=N, half=half, ms=round(ms, 3), Gpairs_per_s=round(N * N / ms / 1e6, 1))

    if want("cls"):
        out = torch.randn(B, 1600, 4097, device=dev, dtype=dt)
        cls = out.permute(0, 2, 1).reshape(B, 4097, 40, 40)[:, :-1]
        alg = B * 1600 * 4096 * es
        ms = timeit(lambda: ops.cls_to_flow_refine(cls), args.iters, flush)
        emit(kernel="cls_to_flow_refine", batch=B, dtype=args.dtype, ms=round(ms, 4), alg_MB=round(alg / 1e6, 2),
             GBps=round(alg / ms / 1e6, 1), frac_hbm=round(alg / ms / 1e6 / PEAKS["hbm_gbs"], 4))

    if want("tiny"):
        Bt = 8
        f0 = torch.randn(Bt, 64, 60, 80, device=dev); f1 = torch.randn(Bt, 64, 60, 80, device=dev)
        flops = Bt * 2 * 4800 * 4800 * 64
        ms = timeit(lambda: ops.tiny_corr_posembed(f0, f1), max(3, args.iters // 4), flush)
        emit(kernel="tiny_corr_posembed", batch=Bt, ms=round(ms, 3), alg_GF=round(flops / 1e9, 2), TFLOPs=round(flops / ms / 1e9, 2))

    if want("dwconv"):
        # ConvRefiner bodies at the 560 -> 864 config: (padded channels, side) of every refiner visit
        for name, Cp, side in (("s16", 1144, 40), ("s8", 1144, 70), ("s4", 576, 140), ("s2", 144, 280), ("s1", 24, 560),
                               ("s8u", 1144, 108), ("s4u", 576, 216), ("s2u", 144, 432), ("s1u", 24, 864)):
            x = torch.randn(B, side, side, Cp, device=dev, dtype=dt).permute(0, 3, 1, 2)
            wt = torch.randn(25, Cp, device=dev) * 0.2
            bs = torch.randn(Cp, device=dev)
            alg = 2 * x.numel() * x.element_size()
            macs = 25 * x.numel()
            ms = timeit(lambda: ops.dwconv5x5_bias_act(x, wt, bs, True), args.iters, flush)
            emit(kernel="dwconv5x5", visit=name, C=Cp, side=side, batch=B, dtype=args.dtype, ms=round(ms, 4),
                 alg_MB=round(alg / 1e6, 1), GBps=round(alg / ms / 1e6, 1), frac_hbm=round(alg / ms / 1e6 / PEAKS["hbm_gbs"], 4),
                 fp32_TFLOPs=round(2 * macs / ms / 1e9, 2))

    if want("fused24"):
        # the whole finest-scale refiner block (depthwise 5x5 + BN + ReLU + 1x1, 24 channels) as one launch
        for name, side in (("s1", 560), ("s1u", 864)):
            x = torch.randn(B, side, side, 24, device=dev, dtype=dt).permute(0, 3, 1, 2)
            wt, bs = torch.randn(25, 24, device=dev) * 0.2, torch.randn(24, device=dev)
            pw, pb = torch.randn(24, 24, device=dev) * 0.2, torch.randn(24, device=dev)
            alg = 2 * x.numel() * x.element_size()
            ms = timeit(lambda: ops.dwconv5x5_pointwise(x, wt, bs, pw, pb, relu=True), args.iters, flush)
            emit(kernel="dwconv5x5_pointwise", visit=name, side=side, batch=B, dtype=args.dtype, ms=round(ms, 4),
                 alg_MB=round(alg / 1e6, 1), GBps=round(alg / ms / 1e6, 1), frac_hbm=round(alg / ms / 1e6 / PEAKS["hbm_gbs"], 4))

    if want("narrow"):
        for name, Cin, Cout, side in (("pw24 s1u", 24, 24, 864), ("pw24 s1", 24, 24, 560), ("proj1 864", 64, 9, 864)):
            x = torch.randn(B, side, side, Cin, device=dev, dtype=dt).permute(0, 3, 1, 2)
            wt = torch.randn(Cout, Cin, device=dev) * 0.2
            bs = torch.randn(Cout, device=dev)
            alg = B * side * side * (Cin + Cout) * x.element_size()
            ms = timeit(lambda: ops.pointwise_narrow(x, wt, bs), args.iters, flush)
            emit(kernel="pointwise_narrow", case=name, batch=B, ms=round(ms, 4), alg_MB=round(alg / 1e6, 1),
                 GBps=round(alg / ms / 1e6, 1), frac_hbm=round(alg / ms / 1e6 / PEAKS["hbm_gbs"], 4))
    if want("outconv"):
        for name, Cp, side in (("s1u", 24, 864), ("s2u", 144, 432), ("s4u", 576, 216), ("s8u", 1144, 108)):
            x = torch.randn(B, side, side, Cp, device=dev, dtype=dt).permute(0, 3, 1, 2)
            wt = torch.randn(3, Cp, device=dev) * 0.1
            bs = torch.randn(3, device=dev)
            alg = B * side * side * (Cp * x.element_size() + 16)
            ms = timeit(lambda: ops.refiner_out_conv(x, wt, bs), args.iters, flush)
            emit(kernel="refiner_out_conv", case=name, batch=B, ms=round(ms, 4), alg_MB=round(alg / 1e6, 1),
                 GBps=round(alg / ms / 1e6, 1), frac_hbm=round(alg / ms / 1e6 / PEAKS["hbm_gbs"], 4))
    if want("stream"):
        x = torch.randn(B, 1601, 1024, device=dev, dtype=dt)
        g1, b1 = torch.ones(1024, device=dev, dtype=dt), torch.zeros(1024, device=dev, dtype=dt)
        alg = 2 * x.numel() * x.element_size()
        ms = timeit(lambda: ops.layer_norm(x, g1, b1, 1e-6), args.iters, flush)
        emit(kernel="layernorm", rows=B * 1601, D=1024, ms=round(ms, 4), GBps=round(alg / ms / 1e6, 1), frac_hbm=round(alg / ms / 1e6 / PEAKS["hbm_gbs"], 4))
        for C, side in ((64, 864), (128, 432), (256, 216)):
            y = torch.randn(B, side, side, C, device=dev, dtype=dt).permute(0, 3, 1, 2)
            alg = y.numel() * y.element_size() * 5 // 4
            ms = timeit(lambda: ops.maxpool2x2(y), args.iters, flush)
            emit(kernel="maxpool2x2", C=C, side=side, ms=round(ms, 4), GBps=round(alg / ms / 1e6, 1), frac_hbm=round(alg / ms / 1e6 / PEAKS["hbm_gbs"], 4))


if __name__ == "__main__":
    main()
