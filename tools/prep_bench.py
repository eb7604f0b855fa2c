"""roma_prepare_images alone: python tools/prep_bench.py  -> one JSON line per case (CUDA events, L2 flushed between launches)."""
import json
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from roma_b200 import ops  # noqa: E402

dev = torch.device("cuda:0")
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for B, H, W, h, w in [(8, 864, 864, 560, 560), (8, 864, 864, 864, 864), (8, 480, 640, 560, 560), (4, 2240, 2240, 560, 560)]:
    img = torch.randint(0, 256, (B, 3, H, W), dtype=torch.uint8, device=dev)
    for _ in range(3):
        ops.prepare_images(img, (h, w))
    ts = []
    for _ in range(10):
        flush.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); ops.prepare_images(img, (h, w)); e.record(); torch.cuda.synchronize()
        ts.append(s.elapsed_time(e))
    ms = sorted(ts)[len(ts) // 2]
    nbytes = B * 3 * (H * W + 4 * h * w)
    print(json.dumps({"case": [B, H, W, h, w], "ms": round(ms, 4), "alg_MB": round(nbytes / 1e6, 1), "GBps": round(nbytes / ms / 1e6, 1)}))

# PIL-exact path: interleaved 8-bit images as decoded, vs Pillow itself on one host thread
import time  # noqa: E402

import numpy as np  # noqa: E402
from PIL import Image  # noqa: E402

for B, H, W, h, w in [(8, 864, 864, 560, 560), (8, 768, 1024, 560, 560), (8, 768, 1024, 864, 864), (2, 1200, 1600, 560, 560)]:
    host = np.random.default_rng(0).integers(0, 256, size=(B, H, W, 3), dtype=np.uint8)
    img = torch.from_numpy(host).to(dev)
    for _ in range(3):
        ops.prepare_images_pil(img, (h, w), channels_last=True)
    ts = []
    for _ in range(10):
        flush.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); ops.prepare_images_pil(img, (h, w), channels_last=True); e.record(); torch.cuda.synchronize()
        ts.append(s.elapsed_time(e))
    ms = sorted(ts)[len(ts) // 2]
    t0 = time.perf_counter()
    for i in range(B):
        np.asarray(Image.fromarray(host[i]).resize((w, h), resample=Image.BICUBIC), dtype=np.float32)
    pil_ms = (time.perf_counter() - t0) * 1e3
    nbytes = B * 3 * (H * W + 4 * h * w)
    print(json.dumps({"case": ["pil", B, H, W, h, w], "ms": round(ms, 4), "alg_MB": round(nbytes / 1e6, 1),
                      "GBps": round(nbytes / ms / 1e6, 1), "pillow_host_ms": round(pil_ms, 2)}))
