"""Tiny-RoMa end-to-end golden fixture: the UNMODIFIED reference `TinyRoMa.match` (romatch/models/tiny.py) on the
CPU, driven with the seeded XFeat-architecture trunk of `roma_b200.host.tiny.XFeatBackbone` (the real XFeat is a
torch.hub download and is not part of the reference tree).

Run in the build container only (needs /root/reference):

    python tests/golden/make_golden_tiny.py

Writes tests/golden/tiny_e2e.npz: the two images, a digest of the weights, and (warp, certainty) of `match`
plus the stride-8 / stride-4 correspondences of `forward`.
"""
from __future__ import annotations

import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
from tests.golden.make_golden import import_reference, save  # noqa: E402
from tests.golden.make_golden_e2e import synthetic_image  # noqa: E402


def main():
    torch.set_num_threads(8)
    import_reference()
    from romatch.models.tiny import TinyRoMa as RefTiny
    import roma_b200
    from roma_b200.host.tiny import XFeatBackbone, random_init_tiny

    weights, trunk_sd = random_init_tiny(seed=0)
    trunk = XFeatBackbone()
    trunk.load_state_dict(trunk_sd)
    ref = RefTiny(xfeat=trunk)
    ref.load_state_dict(weights)
    ref.train(False)

    g = torch.Generator().manual_seed(77)
    im_A = synthetic_image(g, 200, 232)          # not multiples of 32: preprocess_tensor resizes to 192 x 224
    im_B = synthetic_image(g, 200, 232)
    tA = torch.from_numpy(im_A).permute(2, 0, 1)[None].float() / 255
    tB = torch.from_numpy(im_B).permute(2, 0, 1)[None].float() / 255
    with torch.inference_mode():
        warp, cert = ref.match(tA, tB, batched=True)
        corr = ref.forward({"im_A": tA, "im_B": tB})
    print("warp", tuple(warp.shape), "cert mean", float(cert.mean()), "flow8 std", float(corr[8]["flow"].std()))
    save("tiny_e2e", im_A=im_A, im_B=im_B,
         digest=np.frombuffer(roma_b200.state_digest({**weights, **{"xfeat." + k: v for k, v in trunk_sd.items()}}).encode(),
                              dtype=np.uint8),
         warp=warp[:, ::2, ::2].contiguous(), certainty=cert[:, ::2, ::2].contiguous(),
         flow8=corr[8]["flow"], cert8=corr[8]["certainty"], flow4=corr[4]["flow"], cert4=corr[4]["certainty"])


if __name__ == "__main__":
    main()
