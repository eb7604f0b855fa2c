"""Which seam of patch() moves the output of the real reference model?  (GPU box; needs baseline/_ref)
python tools/patch_diag.py"""
import sys
import types
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import roma_b200  # noqa: E402
import importlib  # noqa: E402
from roma_b200 import ops  # noqa: E402
P = importlib.import_module('roma_b200.patch')
from oracle import reference as R  # noqa: E402

torch.backends.cuda.matmul.allow_tf32 = False
torch.backends.cudnn.allow_tf32 = False
dev = torch.device("cuda:0")
w, dino = roma_b200.random_init_weights(seed=0, sensitised=True)
g = torch.Generator().manual_seed(1)
import torch.nn.functional as F  # noqa: E402
im = torch.rand(2, 1, 3, 864, 864, generator=g).to(dev)
fA, fB = im[0] * 2 - 1, im[1] * 2 - 1
dn = lambda t: F.interpolate(t, size=(560, 560), mode="bicubic", antialias=True, align_corners=False)
cA, cB = dn(fA), dn(fB)


def run(model):
    return R.reference_forward_batched(model, cA, cB, fA, fB)


def stats(a, b):
    d = (a - b).abs().flatten()
    return f"max {d.max().item():.2e} p99.9 {d.float().kthvalue(int(d.numel() * 0.999)).values.item():.2e} mean {d.mean().item():.2e}"


ref = R.reference_roma_outdoor(dev, w, dino, 560, 864, amp_dtype=torch.float32)
import romatch.models.matcher as M  # noqa: E402
base = run(ref)
again = run(ref)
print("unpatched twice:           warp", stats(base[0], again[0]), "| logits", stats(base[2], again[2]))

orig = dict(lc=M.local_correlation, cls=M.cls_to_flow_refine)
# 1. GP only
for gp in ref.decoder.gps.values():
    gp.forward = types.MethodType(P._gp_forward, gp)
o = run(ref)
print("GP.forward only:           warp", stats(base[0], o[0]), "| logits", stats(base[2], o[2]))
for gp in ref.decoder.gps.values():
    del gp.forward
# 2. cls_to_flow_refine only
M.cls_to_flow_refine = P._cls_to_flow_refine
o = run(ref)
print("cls_to_flow_refine only:   warp", stats(base[0], o[0]), "| logits", stats(base[2], o[2]))
M.cls_to_flow_refine = orig["cls"]
# 3. module-level local_correlation only
M.local_correlation = ops.local_correlation
o = run(ref)
print("local_correlation only:    warp", stats(base[0], o[0]), "| logits", stats(base[2], o[2]))
M.local_correlation = orig["lc"]
# 4. ConvRefiner.forward only, one scale at a time
for name, rf in ref.decoder.conv_refiner.items():
    rf.forward = types.MethodType(P._refiner_forward, rf)
    o = run(ref)
    print(f"ConvRefiner[{name}].forward only: warp", stats(base[0], o[0]), "| logits", stats(base[2], o[2]))
    del rf.forward
