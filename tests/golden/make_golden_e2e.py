"""End-to-end golden fixtures: the UNMODIFIED reference `RegressionMatcher.match` on the CPU, small resolution.

Run in the build container only (needs /root/reference):

    python tests/golden/make_golden_e2e.py

For each of two seeded random inits (PyTorch default; "sensitised") it
  1. builds the weights with `roma_b200.random_init_weights(seed)` (key set identical to the reference's),
  2. loads them into the reference `romatch.models.model_zoo.roma_models.roma_model` on the CPU (fp32),
  3. writes two synthetic RGB images to PNG files and calls the reference `model.match(pathA, pathB, device="cpu")`,
  4. stores images, the weight digest, and a 2x-subsampled copy of (warp, certainty) in tests/golden/e2e_<tag>.npz.
`tests/test_e2e_oracle.py` checks oracle/roma_model_oracle.py against these on the CPU; the GPU test checks the
CUDA path against the oracle on the same inputs.
"""
from __future__ import annotations

import sys
import tempfile
from pathlib import Path

import numpy as np
import torch
import torch.nn.functional as F
from PIL import Image

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
from tests.golden.make_golden import import_reference, save  # noqa: E402

COARSE, UPSAMPLE = (112, 126), (168, 196)       # non-square on purpose; multiples of 14


def synthetic_image(g: torch.Generator, h: int, w: int) -> np.ndarray:
    """Smooth random field + texture, uint8 RGB (compresses well, has structure at several scales)."""
    low = torch.rand(1, 3, 6, 7, generator=g)
    mid = torch.rand(1, 3, 24, 28, generator=g)
    im = 0.6 * F.interpolate(low, size=(h, w), mode="bicubic", align_corners=False) \
        + 0.4 * F.interpolate(mid, size=(h, w), mode="bilinear", align_corners=False)
    return (im.clamp(0, 1)[0].permute(1, 2, 0) * 255).round().to(torch.uint8).numpy()


def main():
    torch.set_num_threads(8)
    import_reference()
    from romatch.models.model_zoo.roma_models import roma_model as ref_roma_model
    import roma_b200

    g = torch.Generator().manual_seed(2024)
    im_A = synthetic_image(g, 150, 200)
    im_B = synthetic_image(g, 180, 170)
    tmp = Path(tempfile.mkdtemp(prefix="roma_e2e_"))
    Image.fromarray(im_A).save(tmp / "A.png")
    Image.fromarray(im_B).save(tmp / "B.png")

    for tag, sens in (("default", False), ("sensitised", True)):
        weights, dinov2 = roma_b200.random_init_weights(seed=0, sensitised=sens)
        ref = ref_roma_model(COARSE, True, device="cpu", weights=weights, dinov2_weights=dinov2, amp_dtype=torch.float32)
        ref.upsample_res = UPSAMPLE
        warp, cert = ref.match(str(tmp / "A.png"), str(tmp / "B.png"), device="cpu")
        clamped = float(((warp[..., 2:].abs() >= 1).any(-1) | (warp[..., :2].abs() >= 1).any(-1)).float().mean())
        print(tag, "warp", tuple(warp.shape), "cert", tuple(cert.shape), "clamped frac", clamped,
              "cert mean", float(cert.mean()))
        # asymmetric, single pass, no attenuation: the other branches of match()
        ref.symmetric, ref.upsample_preds, ref.attenuate_cert = False, False, False
        warp1, cert1 = ref.match(str(tmp / "A.png"), str(tmp / "B.png"), device="cpu")
        save(f"e2e_{tag}", im_A=im_A, im_B=im_B, coarse=np.array(COARSE), upsample=np.array(UPSAMPLE),
             digest=np.frombuffer(roma_b200.state_digest(weights).encode(), dtype=np.uint8),
             warp=warp[::2, ::2].contiguous(), certainty=cert[::2, ::2].contiguous(),
             warp_single=warp1[::2, ::2].contiguous(), certainty_single=cert1[::2, ::2].contiguous())


if __name__ == "__main__":
    main()
