"""BASELINE.json configs[0] at its stated size: the UNMODIFIED reference `RegressionMatcher.match` on
assets/sacre_coeur_A.jpg / sacre_coeur_B.jpg, coarse 560 x 560 -> upsample 864 x 864, CPU fp32, seeded random init
(PyTorch default and "sensitised").

Run in the build container only (needs /root/reference):

    python tests/golden/make_golden_config1.py

Stores, per init, the 4x-subsampled (warp, certainty) of the reference's symmetric two-pass match (216 x 432 samples of
the 864 x 1728 field) plus the weight digest; the two DECODED images (uint8, as PIL decodes the JPEGs in this image) go into
config1_images.npz once, so the GPU test does not depend on the JPEG decoder or on the asset files being on the box.
"""
from __future__ import annotations

import sys
import time
from pathlib import Path

import numpy as np
import torch
from PIL import Image

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
from tests.golden.make_golden import REF, import_reference, save  # noqa: E402

COARSE, UPSAMPLE, SUB = (560, 560), (864, 864), 4


def _unbatched_inv():
    """torch.linalg.inv on a CPU batch dead-locks inside oneMKL for n >~ 100 in this torch build: run the
    SAME LAPACK routine one matrix at a time (identical arithmetic per element).  The reference source is untouched."""
    inv = torch.linalg.inv

    def one_at_a_time(A, *a, **k):
        if A.dim() == 3 and A.device.type == "cpu" and A.shape[0] > 1:
            return torch.stack([inv(x, *a, **k) for x in A])
        return inv(A, *a, **k)
    torch.linalg.inv = one_at_a_time


def main():
    torch.set_num_threads(8)
    _unbatched_inv()
    import_reference()
    from romatch.models.model_zoo.roma_models import roma_model as ref_roma_model
    import roma_b200

    pA, pB = REF / "assets" / "sacre_coeur_A.jpg", REF / "assets" / "sacre_coeur_B.jpg"
    im_A, im_B = np.array(Image.open(pA).convert("RGB")), np.array(Image.open(pB).convert("RGB"))
    save("config1_images", im_A=im_A, im_B=im_B)
    for tag, sens in (("default", False), ("sensitised", True)):
        weights, dinov2 = roma_b200.random_init_weights(seed=0, sensitised=sens)
        ref = ref_roma_model(COARSE, True, device="cpu", weights=weights, dinov2_weights=dinov2, amp_dtype=torch.float32)
        ref.upsample_res = UPSAMPLE
        t0 = time.time()
        warp, cert = ref.match(str(pA), str(pB), device="cpu")
        print(tag, "reference match on the CPU:", round(time.time() - t0, 1), "s; warp", tuple(warp.shape),
              "cert mean", float(cert.mean()), "zeroed frac", float((cert == 0).float().mean()))
        save(f"config1_{tag}", coarse=np.array(COARSE), upsample=np.array(UPSAMPLE), sub=SUB,
             digest=np.frombuffer(roma_b200.state_digest(weights).encode(), dtype=np.uint8),
             warp=warp[::SUB, ::SUB].contiguous(), certainty=cert[::SUB, ::SUB].contiguous())


if __name__ == "__main__":
    main()
