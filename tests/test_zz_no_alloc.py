"""roma_match_keypoints_static — the entry point that allocates nothing (scratch in a file-scope device array) —
returns exactly what roma_match_keypoints returns, twice in a row, called straight through the C ABI.

Collected last on purpose (file name): a failure here hides nothing after it.
Integer outputs -> the comparison is torch.equal, the bar every keypoint test in this suite uses."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _case(seed, Na, Nb, H, W, dev):
    g = torch.Generator().manual_seed(seed)
    ys = torch.linspace(-1 + 1 / H, 1 - 1 / H, H)
    xs = torch.linspace(-1 + 1 / W, 1 - 1 / W, W)
    grid = torch.stack((xs[None, :].expand(H, W), ys[:, None].expand(H, W)), dim=-1)
    to = (grid + 0.1 * torch.sin(3 * grid.flip(-1)) + 0.01 * torch.randn(H, W, 2, generator=g)).clamp(-1, 1)
    warp = torch.cat((grid, to), dim=-1).contiguous()
    cert = torch.rand(H, W, generator=g)
    xA = torch.rand(Na, 2, generator=g) * 2 - 1
    xB = torch.rand(Nb, 2, generator=g) * 2 - 1
    return [t.to(dev).contiguous() for t in (xA, xB, warp, cert)]


def _call(lib, static, xA, xB, warp, cert, thresh=0.05):
    dev = xA.device
    Na, Nb = xA.shape[0], xB.shape[0]
    H, Wp, _ = warp.shape
    cap = max(Na, Nb)
    ia = torch.full((cap,), -1, device=dev, dtype=torch.int64)
    ib = torch.full((cap,), -1, device=dev, dtype=torch.int64)
    count = torch.zeros(1, device=dev, dtype=torch.int32)
    st = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    p = lambda t: C.c_void_p(t.data_ptr())
    if static:
        rc = lib.roma_match_keypoints_static(p(xA), Na, p(xB), Nb, p(warp), p(cert), H, Wp, thresh, p(ia), p(ib), cap, p(count), st)
    else:
        ws = torch.empty(lib.roma_match_keypoints_workspace_bytes(Na, Nb), device=dev, dtype=torch.uint8)
        rc = lib.roma_match_keypoints(p(xA), Na, p(xB), Nb, p(warp), p(cert), H, Wp, thresh, p(ia), p(ib), cap, p(count),
                                      p(ws), ws.numel(), st)
    torch.cuda.synchronize(dev)
    return rc, ia, ib, int(count.item())


def test_static_scratch_entry_point_matches_the_workspace_one_twice(cuda):
    from roma_b200 import _lib
    lib = _lib.load()
    assert lib.roma_match_keypoints_static_max_points() == 16384
    with torch.cuda.device(cuda):
        # two calls in a row on DIFFERENT sizes: the second finds the first one's leftovers in the shared array
        for seed, Na, Nb in ((1, 3000, 2500), (2, 1200, 1700), (3, 16384, 16384)):
            args = _case(seed, Na, Nb, 48, 96, cuda)
            rc0, ia0, ib0, n0 = _call(lib, False, *args)
            rc1, ia1, ib1, n1 = _call(lib, True, *args)
            rc2, ia2, ib2, n2 = _call(lib, True, *args)
            assert rc0 == 0 and rc1 == 0 and rc2 == 0, lib.roma_last_error_string()
            assert n0 > 0 and n1 == n0 and n2 == n0
            assert torch.equal(ia1, ia0) and torch.equal(ib1, ib0)
            assert torch.equal(ia2, ia0) and torch.equal(ib2, ib0)


def test_static_scratch_entry_point_rejects_sizes_beyond_its_limit(cuda):
    from roma_b200 import _lib
    lib = _lib.load()
    with torch.cuda.device(cuda):
        args = _case(4, 16385, 100, 16, 16, cuda)
        rc, _, _, _ = _call(lib, True, *args)
        assert rc == -2                                                         # ROMA_ERR_UNSUPPORTED
        assert b"16384" in lib.roma_last_error_string()
        args = _case(5, 100, 20000, 16, 16, cuda)
        assert _call(lib, True, *args)[0] == -2
