"""BASELINE config 5 on one GPU: tiny_roma_v1_outdoor (XFeat stand-in trunk, seeded random init), synthetic
480 x 640 pairs, `match` on a batch + `sample(5000)` per pair.  python tools/tiny_bench.py [pairs_per_step=32] [steps=5]

Prints one JSON line: pairs/s (CUDA events, after 3 warm-up steps) and the live per-kernel timings of the
own kernels on the path (fused correlation + soft-argmax, gathers, kde).  Config 5 proper is 256 pairs sharded
32 per GPU over 8 GPUs with no collective (pairs are independent), i.e. 8 replicas of this run."""
import json
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from roma_b200 import ops  # noqa: E402
from roma_b200.host.tiny import TinyRoMa, XFeatBackbone, random_init_tiny  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
dev = torch.device("cuda:0")
weights, trunk_sd = random_init_tiny(seed=0)
trunk = XFeatBackbone()
trunk.load_state_dict(trunk_sd)
model = TinyRoMa(trunk)
model.load_state_dict(weights)
model = model.to(dev).train(False)
g = torch.Generator().manual_seed(4321)
A = torch.rand(B, 3, 480, 640, generator=g).to(dev)
Bm = torch.rand(B, 3, 480, 640, generator=g).to(dev)


def step():
    warp, cert = model.match(A, Bm, batched=True)
    for i in range(B):
        model.sample(warp[i], cert[i], num=5000)


for _ in range(3):
    step()
torch.cuda.synchronize()
s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
with ops.KernelTimer() as kt:
    s.record()
    for _ in range(steps):
        step()
    e.record()
    torch.cuda.synchronize()
ms = s.elapsed_time(e) / steps
rows = []
for name, recs in kt.summary().items():
    tot = sum(m for _, m in recs)
    rows.append({"kernel": name, "launches_per_step": len(recs) / steps, "ms_per_step": round(tot / steps, 3),
                 "avg_ms": round(tot / len(recs), 4), "dims": recs[0][0]})
rows.sort(key=lambda r: -r["ms_per_step"])
print(json.dumps({"metric": "image_pairs_per_s_tiny_roma_480x640", "value": round(B / ms * 1e3, 1), "unit": "pairs/s",
                  "pairs_per_step": B, "steps": steps, "ms_per_step": round(ms, 3), "dtype": "f32 trunk, f16 kde",
                  "kernels": rows}))
