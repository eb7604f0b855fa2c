"""config 1 (photographs, random init): how the GP arithmetic moves the ill-conditioned coarse argmax.
Fraction of pixels whose warp differs by > 1e-3 from the reference on the same GPU (fp32), per roma_gp_set_path."""
import sys
import tempfile
from pathlib import Path

import numpy as np
import torch
from PIL import Image

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import roma_b200  # noqa: E402
from roma_b200 import _lib  # noqa: E402
from oracle import reference as R  # noqa: E402

torch.backends.cuda.matmul.allow_tf32 = False
torch.backends.cudnn.allow_tf32 = False
dev = torch.device("cuda:0")
z = np.load(ROOT / "tests/golden/config1_images.npz")
tmp = Path(tempfile.mkdtemp())
Image.fromarray(z["im_A"]).save(tmp / "A.png"); Image.fromarray(z["im_B"]).save(tmp / "B.png")
pA, pB = str(tmp / "A.png"), str(tmp / "B.png")
lib = _lib.load()
for tag in ("default", "sensitised"):
    w, dino = roma_b200.random_init_weights(seed=0, sensitised=tag == "sensitised")
    g = np.load(ROOT / f"tests/golden/config1_{tag}.npz")
    ref = R.reference_roma_outdoor(dev, w, dino, 560, 864, amp_dtype=torch.float32)
    rw, rc = ref.match(pA, pB, device=dev)
    del ref
    torch.cuda.empty_cache()
    gw = torch.from_numpy(g["warp"])
    print(tag, "reference CUDA vs CPU golden: frac>1e-3", ((rw[::4, ::4].cpu() - gw).abs() > 1e-3).float().mean().item())
    model = roma_b200.roma_outdoor(dev, w, dino, coarse_res=560, upsample_res=864, amp_dtype=torch.float32)
    for path, name in ((0, "2-plane split (default)"), (2, "3-plane split"), (1, "fp32 SIMT Cholesky")):
        lib.roma_gp_set_path(path)
        ww, cc = model.match(pA, pB, device="cuda")
        d = (ww - rw).abs()
        print(f"  {tag} gp path {path} [{name}]: vs reference CUDA frac>1e-3 {(d > 1e-3).float().mean().item():.5f} max {d.max().item():.3e} "
              f"mean {d.mean().item():.3e}; vs CPU golden frac {((ww[::4, ::4].cpu() - gw).abs() > 1e-3).float().mean().item():.5f}; "
              f"cert max {(cc - rc).abs().max().item():.2e}")
    lib.roma_gp_set_path(0)
    del model
    torch.cuda.empty_cache()
