"""CPU oracle for the PIL bicubic resize of 8-bit images — TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s CPU-baseline legs may import this module.

Why it exists: the reference resizes every input image with `PIL.Image.resize(size, resample=Image.BICUBIC)`
before ToTensor / Normalize (romatch/utils/utils.py:165-174 `get_tuple_transform_ops`, :223-232 `TupleResize`,
called from romatch/models/matcher.py:628-634 and :667-676).  The arithmetic lives in a THIRD-PARTY dependency
that is not under /root/reference: Pillow (requirements.txt lists it unpinned; this image has Pillow 12.2.0),
`src/libImaging/Resample.c`.  Its published algorithm for 8-bit channels is restated here:

  * per output index, coefficients in double precision: `scale = in / out`, `filterscale = max(scale, 1)`,
    `support = 2 * filterscale` (bicubic), `center = (i + 0.5) * scale`, `xmin = max(int(center - support + 0.5), 0)`,
    `xmax = min(int(center + support + 0.5), in) - xmin`, `w_x = bicubic((x + xmin - center + 0.5) / filterscale)`
    (evaluated as `* (1 / filterscale)`), normalised by their running sum;
  * coefficients quantised to fixed point with 22 fractional bits, rounding half away from zero;
  * horizontal pass first, then vertical, EACH producing 8-bit pixels: `clip8((2^21 + sum u8 * k) >> 22)`;
    a pass is skipped when its dimension does not change.

Pinning: `tests/test_pil_resample_oracle.py` checks this restatement for exact equality with Pillow itself
(the dependency, executed in the test process) over down-, up- and mixed scalings; the CUDA kernel
`roma_prepare_images_pil` is then compared bit for bit with both.
"""
from __future__ import annotations

import math
from typing import Tuple

import numpy as np

PRECISION_BITS = 32 - 8 - 2


def _bicubic(x: float) -> float:
    a = -0.5
    if x < 0.0:
        x = -x
    if x < 1.0:
        return ((a + 2.0) * x - (a + 3.0)) * x * x + 1
    if x < 2.0:
        return (((x - 5) * x + 8) * x - 4) * a
    return 0.0


def bicubic_coeffs_8bpc(in_size: int, out_size: int) -> Tuple[np.ndarray, np.ndarray, int]:
    """(bounds (out, 2) = first input index and tap count, integer coefficients (out, ksize), ksize)."""
    scale = float(in_size) / out_size
    filterscale = max(scale, 1.0)
    support = 2.0 * filterscale
    ksize = int(math.ceil(support)) * 2 + 1
    bounds = np.zeros((out_size, 2), dtype=np.int64)
    kk = np.zeros((out_size, ksize), dtype=np.int64)
    ss = 1.0 / filterscale
    for i in range(out_size):
        center = (i + 0.5) * scale
        xmin = max(int(center - support + 0.5), 0)
        xmax = min(int(center + support + 0.5), in_size) - xmin
        ws, ww = [], 0.0
        for x in range(xmax):
            w = _bicubic((x + xmin - center + 0.5) * ss)
            ws.append(w)
            ww += w
        for x in range(xmax):
            w = ws[x] / ww if ww != 0.0 else ws[x]
            kk[i, x] = int(-0.5 + w * (1 << PRECISION_BITS)) if w < 0 else int(0.5 + w * (1 << PRECISION_BITS))
        bounds[i] = (xmin, xmax)
    return bounds, kk, ksize


def _pass_8bpc(img: np.ndarray, axis: int, out_size: int) -> np.ndarray:
    """One resampling pass along `axis` of an (H, W, C) uint8 array; int64 accumulation of exact integer products."""
    in_size = img.shape[axis]
    bounds, kk, ksize = bicubic_coeffs_8bpc(in_size, out_size)
    src = np.moveaxis(img, axis, 0).astype(np.int64)                     # (in, ...)
    idx = np.clip(bounds[:, :1] + np.arange(ksize)[None], 0, in_size - 1)   # (out, ksize); padded taps have k = 0
    acc = np.full((out_size,) + src.shape[1:], 1 << (PRECISION_BITS - 1), dtype=np.int64)
    for t in range(ksize):
        acc += src[idx[:, t]] * kk[:, t].reshape((-1,) + (1,) * (src.ndim - 1))
    out = np.clip(acc >> PRECISION_BITS, 0, 255).astype(np.uint8)
    return np.moveaxis(out, 0, axis)


def resize_bicubic_u8(img_hwc: np.ndarray, size_hw: Tuple[int, int]) -> np.ndarray:
    """`Image.fromarray(img).resize((w, h), Image.BICUBIC)` for an (H, W, C) uint8 array."""
    assert img_hwc.dtype == np.uint8 and img_hwc.ndim == 3
    h, w = size_hw
    out = img_hwc
    if w != img_hwc.shape[1]:
        out = _pass_8bpc(out, 1, w)
    if h != img_hwc.shape[0]:
        out = _pass_8bpc(out, 0, h)
    return out
