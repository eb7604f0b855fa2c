"""Locate and import the UNMODIFIED reference (`romatch`) — test / benchmark infrastructure, never on the product path.

Only tests/, __graft_entry__.smoke() and bench.py's reference legs may import this module.

Where the reference lives:
  * /root/reference                in the build container (read-only mount; absent on the GPU box), or
  * <repo>/baseline/_ref           the offline install of the same sources
        python -m pip install --no-index --no-build-isolation --no-deps --target baseline/_ref <copy of /root/reference>
    done by `__graft_entry__.build()` when /root/reference is present.  It is git-ignored (not product source,
    never committed) but travels to the GPU box with the gpurun snapshot, so the CUDA <-> CUDA comparisons
    (kde bit-exactness, sampled indices, patch() on a real reference model, the ATen-on-B200 baseline) can run there.
`import romatch` needs `kornia` (romatch/utils/utils.py:10), which is not installed and is not used on the matching
path: an empty stub package is put on sys.path.
"""
from __future__ import annotations

import os
import sys
import tempfile
from pathlib import Path
from typing import Optional

ROOT = Path(__file__).resolve().parent.parent
_imported = None


def locate() -> Optional[Path]:
    cands = [os.environ.get("ROMA_REFERENCE"), "/root/reference", ROOT / "baseline" / "_ref"]
    for c in cands:
        if c and (Path(c) / "romatch" / "__init__.py").exists():
            return Path(c)
    return None


def available() -> bool:
    return locate() is not None


def import_reference():
    """-> the `romatch` package of the unmodified reference; raises FileNotFoundError when it is not on this machine."""
    global _imported
    if _imported is not None:
        return _imported
    where = locate()
    if where is None:
        raise FileNotFoundError("reference not found: neither /root/reference nor baseline/_ref (run __graft_entry__.build() "
                                "in the build container to install it)")
    try:
        import kornia  # noqa: F401
    except Exception:
        stub = Path(tempfile.mkdtemp(prefix="kornia_stub_"))
        (stub / "kornia").mkdir()
        (stub / "kornia" / "__init__.py").write_text("")
        sys.path.insert(0, str(stub))
    sys.path.insert(0, str(where))
    import romatch
    _imported = romatch
    return romatch


def reference_roma_outdoor(device, weights, dinov2_weights, coarse_res=560, upsample_res=864, amp_dtype=None, all_amp=True):
    """The reference's own `roma_outdoor` (model_zoo/__init__.py:30-52) with the given state_dicts.

    all_amp=True also sets `amp_dtype` on decoder / refiners / embedding decoder / VGG, which the reference factory leaves
    at their fp16 default — that is how an all-fp32 or all-bf16 CUDA comparand is obtained."""
    import contextlib
    import io

    import torch
    romatch = import_reference()
    amp_dtype = amp_dtype or torch.float16
    with contextlib.redirect_stdout(io.StringIO()):
        model = romatch.roma_outdoor(device=device, weights=weights, dinov2_weights=dinov2_weights, coarse_res=coarse_res,
                                     upsample_res=upsample_res, amp_dtype=amp_dtype)
    if all_amp and str(device) != "cpu":
        model.decoder.amp_dtype = amp_dtype
        model.decoder.embedding_decoder.amp_dtype = amp_dtype
        for ref in model.decoder.conv_refiner.values():
            ref.amp_dtype = amp_dtype
        model.encoder.amp_dtype = amp_dtype
        model.encoder.cnn.amp_dtype = amp_dtype
    return model.eval()


_inv_patched = False


def unbatched_cpu_inv() -> None:
    """`torch.linalg.inv` on a CPU batch (the reference's GP.forward, matcher.py:261, with b >= 2) dead-locks inside oneMKL for
    n >~ 100 in this torch build.  Run the SAME LAPACK routine one matrix at a time instead — identical
    arithmetic per element; CUDA tensors and single matrices pass straight through.  The reference source is untouched."""
    global _inv_patched
    if _inv_patched:
        return
    import torch
    inv = torch.linalg.inv

    def one_at_a_time(A, *a, **k):
        if A.dim() == 3 and A.device.type == "cpu" and A.shape[0] > 1 and A.shape[-1] >= 64:
            return torch.stack([inv(x, *a, **k) for x in A])
        return inv(A, *a, **k)
    torch.linalg.inv = one_at_a_time
    _inv_patched = True


def reference_forward_batched(model, im_A, im_B, im_A_up, im_B_up):
    """Batched two-pass matching with the UNMODIFIED reference modules.

    `match(batched=True)` cannot take tensors in this revision, so the batch is driven the way SURVEY.md
    §8(d) prescribes: the reference's own `forward_symmetric` for both passes (matcher.py:516-528), and between / after
    them the arithmetic of matcher.py:656-662 and 684-718 written out with torch ops.
    im_A, im_B: normalised (B,3,h,w) at the coarse resolution; im_A_up, im_B_up: at the upsample resolution.
    Returns (warp (B,H,2W,4), certainty (B,H,2W), certainty logits (B,H,2W))."""
    import math

    import torch
    import torch.nn.functional as F
    B = im_A.shape[0]
    H, W = model.upsample_res
    dev = im_A.device
    if dev.type == "cpu":
        unbatched_cpu_inv()
    with torch.no_grad():
        coarse = model.forward_symmetric({"im_A": im_A, "im_B": im_B})
        low = F.interpolate(coarse[16]["certainty"], size=(H, W), mode="bilinear", align_corners=False)
        low = 0.5 * low * (low < 0)
        sf = math.sqrt(H * W / (model.w_resized * model.h_resized))
        fine = model.forward_symmetric({"im_A": im_A_up, "im_B": im_B_up, "corresps": coarse[1]}, upsample=True,
                                       batched=True, scale_factor=sf)
        flow = fine[1]["flow"].permute(0, 2, 3, 1)                       # (2B, H, W, 2)
        logits = fine[1]["certainty"] - low                              # (2B, 1, H, W)
        prob = logits.sigmoid()
        ys = torch.linspace(-1 + 1 / H, 1 - 1 / H, H, device=dev)
        xs = torch.linspace(-1 + 1 / W, 1 - 1 / W, W, device=dev)
        grid = torch.stack((xs[None, :].expand(H, W), ys[:, None].expand(H, W)), dim=-1)[None].expand(B, H, W, 2)
        outside = (flow.abs() > 1).any(dim=-1)
        prob = torch.where(outside[:, None], torch.zeros_like(prob), prob)
        flow = flow.clamp(-1, 1)
        warp = torch.cat((torch.cat((grid, flow[:B]), dim=-1), torch.cat((flow[B:], grid), dim=-1)), dim=2)
        side = lambda t: torch.cat((t[:B, 0], t[B:, 0]), dim=2)
    return warp, side(prob), side(logits)
