"""Identify, on the B200, the exact arithmetic of the reference's CUDA kde chain (romatch/utils/kde.py:4-12):
  (a) which kernels ATen launches (names, grids, blocks) for cdist / pow / div / exp / sum at the sizes sample() uses;
  (b) which candidate arithmetic reproduces the half GEMM of _euclidean_dist bit for bit (d2 = x1_ @ x2_^T);
  (c) which summation order reproduces `scores.sum(dim=-1)` (half in, fp32 accumulate) bit for bit;
  (d) whether the table of the chain after d2 equals ATen's elementwise half ops on every half bit pattern.
Writes gpurun_out/kde_explore.json.  Test infrastructure only."""
from __future__ import annotations

import ctypes as C
import json
import sys
import time
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from oracle import reference as R  # noqa: E402
from roma_b200 import _lib, ops  # noqa: E402

OUT = ROOT / "gpurun_out"
OUT.mkdir(exist_ok=True)
dev = torch.device("cuda:0")
res = {}


def clustered_points(N, seed=0):
    """rows of a synthetic symmetric warp: (xA, yA, xB, yB) in [-1, 1], B = A + smooth offset, with dense clusters."""
    g = torch.Generator().manual_seed(seed)
    a = torch.rand(N, 2, generator=g) * 2 - 1
    k = N // 3
    a[:k] = (torch.randn(k, 2, generator=g) * 0.05 + torch.tensor([0.3, -0.2])).clamp(-1, 1)
    b = (a + 0.1 * torch.sin(3 * a.flip(1)) + 0.002 * torch.randn(N, 2, generator=g)).clamp(-1, 1)
    return torch.cat((a, b), dim=1).to(dev)


def aug(xh):
    n = xh.pow(2).sum(-1, keepdim=True)
    one = torch.ones_like(n)
    return torch.cat((xh.mul(-2), n, one), -1), torch.cat((xh, one, n), -1)


def profile_chain(N):
    from torch.profiler import ProfilerActivity, profile
    x = clustered_points(N)
    romatch = R.import_reference()
    from romatch.utils.kde import kde
    kde(x[:512])
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
        d = kde(x)
        torch.cuda.synchronize()
    path = OUT / f"kde_trace_{N}.json"
    prof.export_chrome_trace(str(path))
    tr = json.loads(path.read_text())
    ks = []
    for ev in tr.get("traceEvents", []):
        if ev.get("cat") == "kernel":
            a = ev.get("args", {})
            ks.append({"name": ev["name"][:160], "grid": a.get("grid"), "block": a.get("block"), "us": ev.get("dur"),
                       "smem": a.get("shared memory"), "regs": a.get("registers per thread")})
    path.unlink()
    return ks


def d2_candidates(N):
    x = clustered_points(N, seed=1)
    xh = x.half()
    x1, x2 = aug(xh)
    # cdist_impl views 2-D inputs as a batch of one, so _euclidean_dist's matmul is a bmm: take that as THE reference,
    # check that it reproduces torch.cdist itself, and record whether the plain 2-D mm agrees with it
    ref = x1[None].matmul(x2[None].mT)[0]
    refb = ref.view(torch.int16)
    out = {}
    out["bmm_form_reproduces_cdist"] = int((ref.clone().clamp_min_(0).sqrt_().view(torch.int16) != torch.cdist(xh, xh).view(torch.int16)).sum().item())
    out["mm2d_vs_bmm"] = int((x1.matmul(x2.mT).view(torch.int16) != refb).sum().item())
    lib = _lib.load()
    xf = x.float().contiguous()
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    for var in range(5):
        buf = torch.empty((N, N), dtype=torch.int16, device=dev)
        rc = lib.roma_kde_debug_d2(C.c_void_p(xf.data_ptr()), N, var, C.c_void_p(buf.data_ptr()), st)
        assert rc == 0, lib.roma_last_error_string()
        torch.cuda.synchronize()
        out[f"kernel_var{var}"] = int((buf != refb).sum().item())
    # pure-torch candidates: exact sum in fp64, then the possible roundings
    x1d, x2d = x1.double(), x2.double()
    exact = x1d @ x2d.mT                                           # 6 exact products, fp64 sum (exact for these magnitudes)
    out["exact_direct"] = int((exact.half().view(torch.int16) != refb).sum().item())
    out["exact_rn32"] = int((exact.float().half().view(torch.int16) != refb).sum().item())
    rz = (exact.view(torch.int64) & ~((1 << 29) - 1)).view(torch.float64).float()
    out["exact_rz32"] = int((rz.half().view(torch.int16) != refb).sum().item())
    # sequential fp32 orders (products are exact in fp32)
    x1f, x2f = x1.float(), x2.float()
    prods = [x1f[:, k:k + 1] * x2f[:, k][None, :] for k in range(6)]

    def seq(order):
        acc = prods[order[0]].clone()
        for k in order[1:]:
            acc = acc + prods[k]
        return acc
    for name, order in {"seq012345": [0, 1, 2, 3, 4, 5], "seq543210": [5, 4, 3, 2, 1, 0], "seq450123": [4, 5, 0, 1, 2, 3]}.items():
        out[name] = int((seq(order).half().view(torch.int16) != refb).sum().item())
    pair = ((prods[0] + prods[1]) + (prods[2] + prods[3])) + (prods[4] + prods[5])
    out["pairwise"] = int((pair.half().view(torch.int16) != refb).sum().item())
    out["fp32_matmul"] = int(((x1f @ x2f.mT).half().view(torch.int16) != refb).sum().item())
    # is the reference GEMM symmetric?  d2(i,j) == d2(j,i)
    out["ref_asymmetric_pairs"] = int((refb != refb.t()).sum().item())
    out["pairs"] = N * N
    # our tcgen05 GEMM on the zero-padded 8-wide vectors
    pad = lambda t: torch.cat((t, torch.zeros(N, 2, dtype=t.dtype, device=dev)), -1).contiguous()
    try:
        tc = ops.tc_gemm_nt(pad(x1)[None], pad(x2)[None])[0]
        out["tcgen05_f16"] = int((tc.half().view(torch.int16) != refb).sum().item())
    except Exception as e:                                         # noqa: BLE001
        out["tcgen05_f16"] = f"error: {e}"
    return out


def aten_sum_emulate(E, vec, bw, bh, split):
    """E (R, N) half -> half, in the order of ATen's reduce kernel (Reduce.cuh) for a contiguous inner-dim sum:
    T = bw * (bh if split else 1) threads per row, `vec` accumulators per thread over aligned vectors, accumulators
    combined left to right, shuffle-down tree over x (offsets bw/2 .. 1), shared-memory tree over y (bh/2 .. 1)."""
    Rr, N = E.shape
    T = bw * (bh if split else 1)
    S = T * vec
    K = (N + S - 1) // S
    Ef = torch.zeros((Rr, K * S), dtype=torch.float32, device=E.device)
    Ef[:, :N] = E.float()
    Ef = Ef.view(Rr, K, T, vec)
    acc = Ef[:, 0].clone()
    for k in range(1, K):
        acc = acc + Ef[:, k]
    v = acc[..., 0]
    for i in range(1, vec):
        v = v + acc[..., i]
    v = v.view(Rr, (bh if split else 1), bw)
    off = bw // 2
    while off > 0:
        v = v[..., :off] + v[..., off:2 * off]
        off //= 2
    v = v[..., 0]
    off = v.shape[1] // 2
    while off > 0:
        v = v[:, :off] + v[:, off:2 * off]
        off //= 2
    return v[:, 0].half()


def sum_candidates(N, rows=2048):
    x = clustered_points(N, seed=2)
    xh = x.half()
    E = (-torch.cdist(xh[:rows], xh) ** 2 / (2 * 0.1 ** 2)).exp()          # (rows, N) half, contiguous
    ref = E.sum(dim=-1)
    out = {"rows": rows, "N": N}
    for vec in (4, 8, 2):
        for (bw, bh, split) in ((32, 16, True), (32, 16, False), (32, 8, True), (32, 4, True), (32, 32, True), (64, 8, True),
                                (128, 4, True), (512, 1, True)):
            if bw > 32:
                continue
            got = aten_sum_emulate(E, vec, bw, bh, split)
            out[f"vec{vec}_bw{bw}_bh{bh}_{'split' if split else 'warp'}"] = int((got.view(torch.int16) != ref.view(torch.int16)).sum().item())
    out["fp64_exact"] = int((E.double().sum(-1).half().view(torch.int16) != ref.view(torch.int16)).sum().item())
    out["fp32_torch_sum"] = int((E.float().sum(-1).half().view(torch.int16) != ref.view(torch.int16)).sum().item())
    # the full-matrix sum goes through with_32bit_indexing when N*N*2 >= 2^31: check that rows computed in a slice agree
    return out


def full_vs_slice(N):
    """kde(x) over the full N x N matrix vs the same rows computed from a row slice (is the result row-local?)."""
    x = clustered_points(N, seed=3)
    from romatch.utils.kde import kde
    full = kde(x)
    xh = x.half()
    rows = 4096
    outs = {}
    for r0 in (0, N // 2 - 2048, N - rows):
        E = (-torch.cdist(xh[r0:r0 + rows], xh) ** 2 / (2 * 0.1 ** 2)).exp()
        part = E.sum(-1)
        outs[f"rows_{r0}"] = int((part.view(torch.int16) != full[r0:r0 + rows].view(torch.int16)).sum().item())
    return outs


def lut_check():
    lib = _lib.load()
    lut = torch.empty(32768, dtype=torch.float32, device=dev)
    rc = lib.roma_kde_debug_lut(C.c_void_p(lut.data_ptr()), 0.1, C.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0
    bits = torch.arange(0, 0x7C01, dtype=torch.int32, device=dev).to(torch.int16)
    d2 = bits.view(torch.float16)
    r = d2.clone().clamp_min_(0).sqrt_()
    e = (-r ** 2 / (2 * 0.1 ** 2)).exp()
    mism = int((lut[:0x7C01].half().view(torch.int16) != e.view(torch.int16)).sum().item())
    nz = int((e != 0).sum().item())
    last = int(torch.nonzero(e != 0).max().item())
    return {"mismatches": mism, "nonzero_entries": nz, "last_nonzero_bits": last}


def current_modes(N):
    """our kde (every mode) vs the reference chain, full size."""
    x = clustered_points(N, seed=4)
    from romatch.utils.kde import kde
    ref = kde(x)
    lib = _lib.load()
    out = {}
    for mode in (0, 1, 2, 3):
        lib.roma_kde_set_direct(mode)
        got = ops.kde(x, half=True)
        out[f"mode{mode}"] = int((got.view(torch.int16) != ref.view(torch.int16)).sum().item())
    lib.roma_kde_set_direct(0)
    return out


def sum_only():
    """all rows, so that wrong summation orders show (~0.1 % of the rows move by one half ulp)"""
    for N in (4000, 8160, 8192, 20000):
        res[f"sumall_{N}"] = sum_candidates(N, rows=N)
        print("sum(all rows)", N, {k: v for k, v in res[f"sumall_{N}"].items()})


def main():
    if "--sum-only" in sys.argv:
        return sum_only()
    torch.manual_seed(0)
    print("device", torch.cuda.get_device_name(0), "torch", torch.__version__, "cublas", torch.backends.cuda.preferred_blas_library())
    res["allow_fp16_reduced_precision_reduction"] = torch.backends.cuda.matmul.allow_fp16_reduced_precision_reduction
    for N in (4096, 20000, 40000):
        res[f"kernels_{N}"] = profile_chain(N)
        print(N, json.dumps(res[f"kernels_{N}"], indent=0)[:3000])
    res["lut"] = lut_check()
    print("lut", res["lut"])
    for N in (2048, 8192):
        res[f"d2_{N}"] = d2_candidates(N)
        print("d2", N, res[f"d2_{N}"])
    for N in (4096, 8000, 8192, 20000, 40000):
        res[f"sum_{N}"] = sum_candidates(N)
        print("sum", N, res[f"sum_{N}"])
    res["full_vs_slice_40000"] = full_vs_slice(40000)
    print("full_vs_slice", res["full_vs_slice_40000"])
    for N in (20000, 40000):
        res[f"modes_{N}"] = current_modes(N)
        print("modes", N, res[f"modes_{N}"])
    (OUT / "kde_explore.json").write_text(json.dumps(res, indent=1))
    sum_only()


if __name__ == "__main__":
    t0 = time.time()
    try:
        main()
    finally:
        (OUT / ("kde_explore_sum.json" if "--sum-only" in sys.argv else "kde_explore.json")).write_text(json.dumps(res, indent=1))
        print("elapsed", time.time() - t0)
