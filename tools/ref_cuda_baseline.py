"""The UNMODIFIED reference (baseline/_ref) on the B200 through ATen: pairs/s of the same workload as bench.py
(8 synthetic pairs, 560 -> 864, forward_symmetric x2 + epilogue + sample(10000) per pair) — the "same GPU" comparand.
Usage: python tools/ref_cuda_baseline.py [--dtype float16|bfloat16|float32] [--batch 8] [--steps 3]
Prints one JSON line; test / benchmark infrastructure only."""
from __future__ import annotations

import argparse
import json
import sys
import time
from pathlib import Path

import torch
import torch.nn.functional as F

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


def run(dtype="float16", batch=8, steps=3, warmup=1, coarse=(560, 560), upsample=(864, 864), num=10000, per_module=True):
    import roma_b200
    from bench import synthetic_pairs
    from oracle import reference as R
    dev = torch.device("cuda:0")
    amp = getattr(torch, dtype)
    weights, dinov2 = roma_b200.random_init_weights(seed=0)
    model = R.reference_roma_outdoor(dev, weights, dinov2, coarse, upsample, amp_dtype=amp, all_amp=True)
    A8, B8 = synthetic_pairs(batch, upsample, 1234)
    mean = torch.tensor((0.485, 0.456, 0.406), device=dev).view(1, 3, 1, 1)
    std = torch.tensor((0.229, 0.224, 0.225), device=dev).view(1, 3, 1, 1)

    def step():
        fa, fb = (A8.to(dev).float() / 255 - mean) / std, (B8.to(dev).float() / 255 - mean) / std
        rs = lambda t: F.interpolate(t, size=coarse, mode="bicubic", antialias=True, align_corners=False)
        warp, cert, _ = R.reference_forward_batched(model, rs(fa), rs(fb), fa, fb)
        outs = [model.sample(warp[i], cert[i], num=num) for i in range(batch)]
        return outs

    for _ in range(warmup):
        step()
    torch.cuda.synchronize()
    torch.cuda.reset_peak_memory_stats()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / steps
    out = {"impl": "reference_aten_cuda", "dtype": dtype, "pairs_per_step": batch, "ms_per_step": round(dt * 1e3, 2),
           "pairs_per_s": round(batch / dt, 3), "peak_mem_gb": round(torch.cuda.max_memory_allocated() / 2 ** 30, 2)}
    if per_module:
        # sample() alone (kde dominates: N x N half temporaries)
        fa, fb = (A8.to(dev).float() / 255 - mean) / std, (B8.to(dev).float() / 255 - mean) / std
        rs = lambda t: F.interpolate(t, size=coarse, mode="bicubic", antialias=True, align_corners=False)
        warp, cert, _ = R.reference_forward_batched(model, rs(fa), rs(fb), fa, fb)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for i in range(batch):
            model.sample(warp[i], cert[i], num=num)
        torch.cuda.synchronize()
        out["sample_ms_per_pair"] = round((time.perf_counter() - t0) / batch * 1e3, 3)
    return out


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--dtype", default="float16")
    ap.add_argument("--batch", type=int, default=8)
    ap.add_argument("--steps", type=int, default=3)
    a = ap.parse_args()
    print(json.dumps(run(a.dtype, a.batch, a.steps)))
