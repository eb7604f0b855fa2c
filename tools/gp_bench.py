"""GP posterior timing + error: python tools/gp_bench.py [N_side] [batch] [dtype] [paths e.g. 2,3,1] [real]

`real` uses GP inputs captured from the random-init model on CPU (roma_b200/csrc/build/gp_inputs.pt, cond ~1.2e3)."""
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
paths = tuple(int(p) for p in sys.argv[4].split(",")) if len(sys.argv) > 4 else (2, 3, 1)      # 2 = default (3-plane)
real = len(sys.argv) > 5 and sys.argv[5] == "real"
lib = _lib.load()
PLANES_DEFAULT = 1
dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(3)
C = D = 512
if real:
    z = torch.load(ROOT / "roma_b200/csrc/build/gp_inputs.pt")
    x, y, pw, pb = z["x"], z["y"], z["pw"].reshape(D, 2), z["pb"]
    x = x.repeat(B // 2, 1, 1, 1).to(dt); y = y.repeat(B // 2, 1, 1, 1).to(dt); side = 40
else:
    base = O.bilinear_resize(torch.randn(B, C, 6, 6, generator=g), (side, side))
    x = (base + 0.3 * torch.randn(B, C, side, side, generator=g)).to(dt)
    y = (base.roll(1, 3) + 0.3 * torch.randn(B, C, side, side, generator=g)).to(dt)
    pw = torch.randn(D, 2, generator=g) * 0.5; pb = torch.randn(D, generator=g) * 0.5
cl = lambda t: t.to(dev).contiguous(memory_format=torch.channels_last)
xd, yd, pwd, pbd = cl(x), cl(y), pw.to(dev), pb.to(dev)
want = O.gp_posterior(x[:1].float(), y[:1].float(), pw, pb, dtype=torch.float64)
ref32 = O.gp_posterior(x[:1].float(), y[:1].float(), pw, pb)
print(json.dumps({"ref_fp32_LUinv_err_vs_fp64": (ref32 - want).abs().max().item(), "mu_absmax": want.abs().max().item()}))
N = side * side
flops = B * (2.0 * N * N * C * 2 + N ** 3 / 3 + 2.0 * N * N * D + 2.0 * N * N * D)
names = {1: "simt", 2: "tcgen05 split3", 3: "tcgen05 split2"}
for path in paths:
    for lanes, planes in (((1, 1), (1, 0), (0, 0)) if path != 1 else ((1, 0),)):
        lib.roma_gp_set_path(path)
        lib.roma_gp_set_lanes(lanes)
        lib.roma_gp_set_planes(planes)
        for _ in range(2):
            mu = ops.gp_posterior(xd, yd, pwd, pbd)
        torch.cuda.synchronize()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        for _ in range(5):
            mu = ops.gp_posterior(xd, yd, pwd, pbd)
        e.record(); torch.cuda.synchronize()
        ms = s.elapsed_time(e) / 5
        print(json.dumps({"path": names[path], "schedule": ("two-lane" if lanes else "single-stream") + (" + bf16 planes via TMA" if planes else ""), "N": N, "B": B, "dtype": str(dt),
                          "ms": round(ms, 3), "alg_TFLOPs": round(flops / ms / 1e9, 1),
                          "err_vs_fp64": (mu[:1].cpu() - want).abs().max().item()}))
lib.roma_gp_set_path(0)
lib.roma_gp_set_lanes(1)
lib.roma_gp_set_planes(PLANES_DEFAULT)
