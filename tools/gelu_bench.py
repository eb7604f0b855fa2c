"""roma_gelu vs ATen GELU on the DINOv2 MLP shape: python tools/gelu_bench.py"""
import json
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from roma_b200 import ops  # noqa: E402

dev = torch.device("cuda:0")
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for dt in (torch.bfloat16, torch.float16):
    x = torch.randn(25616, 4096, device=dev).to(dt)
    res = {}
    for name, fn in (("roma_gelu", lambda: ops.gelu(x)), ("roma_gelu_inplace", lambda: ops.gelu(x.clone(), inplace=True)),
                     ("aten", lambda: torch.nn.functional.gelu(x))):
        if name == "roma_gelu_inplace":
            continue
        for _ in range(3):
            fn()
        ts = []
        for _ in range(10):
            flush.zero_()
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record(); fn(); e.record(); torch.cuda.synchronize()
            ts.append(s.elapsed_time(e))
        res[name] = sorted(ts)[len(ts) // 2]
    nbytes = 2 * x.numel() * x.element_size()
    print(json.dumps({"dtype": str(dt), "shape": list(x.shape), "alg_MB": nbytes / 1e6,
                      **{k + "_ms": round(v, 4) for k, v in res.items()},
                      **{k + "_GBps": round(nbytes / v / 1e6, 1) for k, v in res.items()}}))
