"""The PIL bicubic resize restatement (oracle/pil_resample_oracle.py) against Pillow itself: exact equality.
Pillow is the third-party dependency that implements the reference's image resize (utils/utils.py:223-232)."""
import numpy as np
import pytest
from PIL import Image

from oracle import pil_resample_oracle as P

CASES = [((48, 64), (35, 40)), ((480, 640), (560, 560)), ((618, 640), (864, 864)), ((100, 75), (333, 290)),
         ((600, 1200), (224, 308)), ((37, 53), (37, 80)), ((64, 48), (20, 48)), ((30, 30), (30, 30)),
         ((1120, 896), (280, 224)), ((5, 7), (16, 3))]


def _img(hw, seed, kind):
    g = np.random.default_rng(seed)
    if kind == "noise":
        return g.integers(0, 256, size=(*hw, 3), dtype=np.uint8)
    if kind == "edges":                                  # saturated blocks: exercises the clipping of over/undershoot
        a = g.integers(0, 2, size=(hw[0] // 4 + 1, hw[1] // 4 + 1, 3)) * 255
        return np.kron(a, np.ones((4, 4, 1)))[: hw[0], : hw[1]].astype(np.uint8)
    yy, xx = np.mgrid[: hw[0], : hw[1]]
    return np.stack([(xx * 3 + yy) % 256, (xx + yy * 2) % 256, (xx * yy) % 256], -1).astype(np.uint8)


@pytest.mark.parametrize("src,dst", CASES)
@pytest.mark.parametrize("kind", ["noise", "edges", "ramp"])
def test_oracle_equals_pillow(src, dst, kind):
    img = _img(src, src[0] * 31 + dst[1], kind)
    want = np.asarray(Image.fromarray(img).resize((dst[1], dst[0]), resample=Image.BICUBIC))
    got = P.resize_bicubic_u8(img, dst)
    assert got.shape == want.shape and got.dtype == np.uint8
    assert np.array_equal(got, want), f"max |diff| {np.abs(got.astype(int) - want.astype(int)).max()}"


def test_coefficients_sum_to_one_and_bounds():
    for n_in, n_out in [(864, 560), (560, 864), (7, 3), (3, 16)]:
        bounds, kk, ksize = P.bicubic_coeffs_8bpc(n_in, n_out)
        assert (bounds[:, 0] >= 0).all() and (bounds[:, 0] + bounds[:, 1] <= n_in).all() and (bounds[:, 1] <= ksize).all()
        assert np.abs(kk.sum(1) - (1 << P.PRECISION_BITS)).max() <= ksize       # rounding of each tap
