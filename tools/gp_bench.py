"""GP posterior timing + error on one shape: python tools/gp_bench.py [N_side] [batch] [dtype] [path]"""
import sys, json
from pathlib import Path
import torch
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from roma_b200 import ops, _lib
from oracle import roma_oracle as O
side = int(sys.argv[1]) if len(sys.argv) > 1 else 40
B = int(sys.argv[2]) if len(sys.argv) > 2 else 16
dt = getattr(torch, sys.argv[3]) if len(sys.argv) > 3 else torch.bfloat16
lib = _lib.load()
dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(3)
C = D = 512
base = O.bilinear_resize(torch.randn(B, C, 6, 6, generator=g), (side, side))
x = (base + 0.3 * torch.randn(B, C, side, side, generator=g)).to(dt)
y = (base.roll(1, 3) + 0.3 * torch.randn(B, C, side, side, generator=g)).to(dt)
pw = torch.randn(D, 2, generator=g) * 0.5; pb = torch.randn(D, generator=g) * 0.5
xd, yd, pwd, pbd = x.to(dev), y.to(dev), pw.to(dev), pb.to(dev)
want = O.gp_posterior(x[:1].float(), y[:1].float(), pw, pb, dtype=torch.float64)
N = side * side
flops = B * (2.0 * N * N * C * 2 + N ** 3 / 3 + 2.0 * N * N * D + 2.0 * N * N * D)
for path in (2, 1):
    lib.roma_gp_set_path(path)
    for _ in range(2):
        mu = ops.gp_posterior(xd, yd, pwd, pbd)
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(5):
        mu = ops.gp_posterior(xd, yd, pwd, pbd)
    e.record(); torch.cuda.synchronize()
    ms = s.elapsed_time(e) / 5
    print(json.dumps({"path": {1: "simt", 2: "tcgen05"}[path], "N": N, "B": B, "dtype": str(dt), "ms": round(ms, 3),
                      "TFLOPs": round(flops / ms / 1e9, 1), "err_vs_fp64": (mu[:1].cpu() - want).abs().max().item()}))
lib.roma_gp_set_path(0)
