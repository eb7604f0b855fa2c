"""One steady-state launch of the kernels changed last in round 1 (symmetric kde, GELU) for `ncu --set full`
(profiles/r1_late_kernels_ncu.md): python tools/late_profile.py"""
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from roma_b200 import _lib, ops  # noqa: E402

dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(3)
centres = torch.rand(40, 4, generator=g) * 1.6 - 0.8
x = (centres[torch.randint(0, 40, (40000,), generator=g)] + 0.05 * torch.randn(40000, 4, generator=g)).clamp(-1, 1).to(dev)
h = torch.randn(25616, 4096, device=dev).to(torch.bfloat16)
lib = _lib.load()
for _ in range(2):
    lib.roma_kde_set_direct(3)
    ops.kde(x, half=True)            # every pair twice
    lib.roma_kde_set_direct(0)
    ops.kde(x, half=True)            # every pair once (default)
    ops.gelu(h)
torch.cuda.synchronize()
