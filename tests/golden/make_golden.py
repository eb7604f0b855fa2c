"""Generate the golden fixtures that pin `oracle/` to the REFERENCE implementation.

Run in the build container only (needs /root/reference, which does not exist on the GPU box):

    python tests/golden/make_golden.py

It imports the unmodified reference (`romatch`, with an empty `kornia` stub package because kornia is not
installed and is not used on this path — SURVEY.md §8c), runs each hot-path function on small seeded
inputs on the CPU (fp32, the reference's own CPU path) and stores inputs + outputs in
`tests/golden/*.npz`.  The fixtures are committed; this script is their provenance.
"""
from __future__ import annotations

import os
import sys
import tempfile
from pathlib import Path

import numpy as np
import torch
import torch.nn.functional as F

REF = Path(os.environ.get("ROMA_REFERENCE", "/root/reference"))
OUT = Path(__file__).resolve().parent


def import_reference():
    stub = Path(tempfile.mkdtemp(prefix="kornia_stub_"))
    (stub / "kornia").mkdir()
    (stub / "kornia" / "__init__.py").write_text("")
    sys.path.insert(0, str(stub))
    sys.path.insert(0, str(REF))
    import romatch  # noqa: F401
    return romatch


def save(name: str, **arrays):
    np.savez_compressed(OUT / f"{name}.npz", **{k: (v.detach().cpu().numpy() if torch.is_tensor(v) else np.asarray(v))
                                                for k, v in arrays.items()})
    print(f"wrote {name}.npz", {k: tuple(v.shape) if hasattr(v, "shape") else () for k, v in arrays.items()})


def smooth_flow(g, B, h, w, amp=0.3, oob=True):
    """pixel grid + smooth displacement + a few wild (out-of-range) vectors."""
    ys = torch.linspace(-1 + 1 / h, 1 - 1 / h, h)
    xs = torch.linspace(-1 + 1 / w, 1 - 1 / w, w)
    grid = torch.stack((xs[None, :].expand(h, w), ys[:, None].expand(h, w)))[None].expand(B, 2, h, w)
    low = torch.randn(B, 2, max(h // 4, 2), max(w // 4, 2), generator=g) * amp
    flow = grid + F.interpolate(low, size=(h, w), mode="bilinear", align_corners=False)
    if oob:
        mask = torch.rand(B, 1, h, w, generator=g) < 0.08
        wild = torch.rand(B, 2, h, w, generator=g) * 2.6 - 1.3
        flow = torch.where(mask, wild, flow)
    return flow.contiguous()


def main():
    torch.manual_seed(0)
    torch.set_num_threads(8)
    import_reference()
    from romatch.utils.local_correlation import local_correlation
    from romatch.utils.kde import kde
    from romatch.utils.utils import cls_to_flow_refine
    from romatch.models.matcher import GP, CosKernel, ConvRefiner, RegressionMatcher
    from romatch.models.tiny import TinyRoMa

    g = torch.Generator().manual_seed(1234)

    # ---- A3 grid_sample + A4 local_correlation (three radii, non-square, out-of-range flows) -------------
    for tag, (B, C, h, w, r) in {"r2": (2, 16, 11, 13, 2), "r3": (2, 24, 9, 12, 3), "r7": (1, 8, 10, 10, 7)}.items():
        f0 = torch.randn(B, C, h, w, generator=g)
        f1 = torch.randn(B, C, h, w, generator=g)
        flow = smooth_flow(g, B, h, w)
        corr = local_correlation(f0, f1, local_radius=r, flow=flow, sample_mode="bilinear")
        xhat = F.grid_sample(f1, flow.permute(0, 2, 3, 1), align_corners=False, mode="bilinear")
        save(f"local_corr_{tag}", f0=f0, f1=f1, flow=flow, r=r, corr=corr, xhat=xhat)
    # aligned (flow=None) branch
    f0 = torch.randn(1, 8, 7, 9, generator=g); f1 = torch.randn(1, 8, 7, 9, generator=g)
    save("local_corr_noflow", f0=f0, f1=f1, r=2, corr=local_correlation(f0, f1, local_radius=2))

    # ---- A5 ConvRefiner input assembly: capture the tensor entering block1 --------------------------------
    for tag, (C, E, r, h, w, sf) in {"corr": (16, 8, 2, 10, 12, 1.5429), "nocorr": (9, 6, None, 12, 10, 1.0)}.items():
        K = (2 * r + 1) ** 2 if r else 0
        dim = 2 * C + E + K
        ref = ConvRefiner(dim, dim, 3, kernel_size=5, dw=True, hidden_blocks=1, displacement_emb="linear",
                          displacement_emb_dim=E, local_corr_radius=r, corr_in_other=True, amp=True).eval()
        captured = {}
        ref.block1.register_forward_pre_hook(lambda m, inp: captured.__setitem__("d", inp[0].detach().clone()))
        x = torch.randn(2, C, h, w, generator=g); y = torch.randn(2, C, h, w, generator=g)
        flow = smooth_flow(g, 2, h, w)
        with torch.no_grad():
            disp, cert = ref(x, y, flow, scale_factor=sf, logits=None)
        save(f"refiner_inputs_{tag}", x=x, y=y, flow=flow, scale_factor=sf, r=(r or 0),
             disp_w=ref.disp_emb.weight, disp_b=ref.disp_emb.bias, d=captured["d"])

    # ---- A6/A7 update + resample: the two reference statements executed on seeded tensors ------------------
    flow = smooth_flow(g, 2, 10, 14, oob=False); cert = torch.randn(2, 1, 10, 14, generator=g)
    delta = torch.randn(2, 3, 10, 14, generator=g) * 3
    ins, Hf, Wf = 8, 80, 112
    disp = ins * torch.stack((delta[:, 0].float() / (4 * Wf), delta[:, 1].float() / (4 * Hf)), dim=1)  # matcher.py:397-398
    flow_u, cert_u = flow + disp, cert + delta[:, 2:3]
    sizes = {"up": (20, 28), "odd": (17, 25), "down": (4, 5)}   # x2, non-integer ratio (40->70-like), downsample (560->108-like)
    save("update_resample", flow=flow, cert=cert, delta=delta, ins=ins, Hf=Hf, Wf=Wf, flow_u=flow_u, cert_u=cert_u,
         **{f"flow_{k}": F.interpolate(flow_u, size=s, mode="bilinear", align_corners=False) for k, s in sizes.items()},
         **{f"cert_{k}": F.interpolate(cert_u, size=s, mode="bilinear", align_corners=False) for k, s in sizes.items()},
         **{f"size_{k}": np.array(s) for k, s in sizes.items()})

    # ---- A1/A2 GP ------------------------------------------------------------------------------------------
    for tag, (b, c, h1, w1, h2, w2, D) in {"sq": (2, 32, 8, 8, 8, 8, 16), "rect": (1, 24, 6, 9, 7, 5, 12)}.items():
        gp = GP(CosKernel, T=0.2, learn_temperature=False, only_attention=False, gp_dim=D, basis="fourier", no_cov=True).eval()
        base = F.interpolate(torch.randn(b, c, 3, 3, generator=g), size=(h1, w1), mode="bilinear")
        x = base + 0.3 * torch.randn(b, c, h1, w1, generator=g)
        y = F.interpolate(base, size=(h2, w2), mode="bilinear") + 0.3 * torch.randn(b, c, h2, w2, generator=g)
        with torch.no_grad():
            mu = gp(x, y)
            Kxy = gp.K(gp.reshape(x), gp.reshape(y))
        save(f"gp_{tag}", x=x, y=y, pos_w=gp.pos_conv.weight, pos_b=gp.pos_conv.bias, mu=mu, Kxy=Kxy)

    # ---- A8 epilogue: the real RegressionMatcher.match code path over stub encoder/decoder -----------------
    class StubDecoder(torch.nn.Module):
        def __init__(self, outs): super().__init__(); self.outs = outs; self.calls = 0
        def forward(self, f1, f2, upsample=False, **kw):
            out = self.outs[self.calls]; self.calls += 1; return out

    class StubEncoder(torch.nn.Module):
        def forward(self, x, upsample=False): return {1: x}

    from PIL import Image
    for tag, (H, W, Hc, Wc, symmetric, att) in {"sym": (18, 22, 12, 10, True, True), "asym": (16, 16, 16, 16, False, False)}.items():
        Bd = 2 if symmetric else 1
        coarse = {16: {"certainty": torch.randn(Bd, 1, 3, 4, generator=g) * 2, "flow": torch.zeros(Bd, 2, 3, 4)},
                  1: {"certainty": torch.randn(Bd, 1, Hc, Wc, generator=g), "flow": smooth_flow(g, Bd, Hc, Wc)}}
        fine = {1: {"certainty": torch.randn(Bd, 1, H, W, generator=g) * 2, "flow": smooth_flow(g, Bd, H, W, amp=0.6)}}
        m = RegressionMatcher(StubEncoder(), StubDecoder([coarse, fine]), h=Hc, w=Wc, upsample_preds=True,
                              symmetric=symmetric, attenuate_cert=att)
        m.upsample_res = (H, W)
        im = Image.fromarray((np.random.RandomState(0).rand(30, 40, 3) * 255).astype(np.uint8))
        warp, certainty = m.match(im, im, device="cpu")
        save(f"epilogue_{tag}", flow=fine[1]["flow"], cert=fine[1]["certainty"], cert16=coarse[16]["certainty"],
             symmetric=int(symmetric), attenuate=int(att), warp=warp, certainty=certainty)

    # ---- A10 kde, A9 sample ----------------------------------------------------------------------------------
    pts = torch.cat((torch.rand(600, 4, generator=g) * 2 - 1, torch.randn(400, 4, generator=g) * 0.05 + 0.3)).clamp(-1, 1)
    save("kde", x=pts, dens_half=kde(pts, std=0.1).float(), dens_f32=kde(pts, std=0.1, half=False),
         dens_down=kde(pts, std=0.1, half=False, down=8))
    H, W = 24, 40
    matcher = RegressionMatcher(StubEncoder(), StubDecoder([]), h=H, w=W)
    warp = torch.cat((smooth_flow(g, 1, H, W, oob=False), smooth_flow(g, 1, H, W, oob=False)), dim=1)[0].permute(1, 2, 0).contiguous()
    cert = torch.rand(H, W, generator=g) * (torch.rand(H, W, generator=g) < 0.7)
    torch.manual_seed(99)
    sm, sc = matcher.sample(warp, cert, num=100)
    save("sample", warp=warp, cert=cert, num=100, seed=99, matches=sm, certainty=sc)

    # ---- A11 match_keypoints ------------------------------------------------------------------------------------
    Hk, Wk = 20, 48
    wk = torch.cat((smooth_flow(g, 1, Hk, Wk, oob=False), smooth_flow(g, 1, Hk, Wk, amp=0.1, oob=False)), dim=1)[0].permute(1, 2, 0).contiguous()
    ck = torch.rand(Hk, Wk, generator=g)
    xA = torch.rand(60, 2, generator=g) * 2 - 1
    xB = torch.rand(70, 2, generator=g) * 2 - 1
    ia, ib = matcher.match_keypoints(xA, xB, wk, ck, return_tuple=True, return_inds=True)
    save("match_keypoints", warp=wk, cert=ck, xA=xA, xB=xB, inds_A=ia, inds_B=ib)

    # ---- (f-1) cls_to_flow_refine -------------------------------------------------------------------------------
    cls = torch.randn(2, 64, 5, 6, generator=g) * 3
    save("cls_to_flow", cls=cls, flow=cls_to_flow_refine(cls))

    # ---- T1/T2 Tiny-RoMa --------------------------------------------------------------------------------------------
    tiny = TinyRoMa.__new__(TinyRoMa)
    torch.nn.Module.__init__(tiny)
    tiny.exact_softmax = False
    tiny.train(False)
    fa = torch.randn(2, 16, 8, 12, generator=g); fb = torch.randn(2, 16, 8, 12, generator=g)
    cv = tiny.corr_volume(fa, fb)
    # pos_embed is called per batch element: the reference's `P_lowres[:,-1] * grid[best_match].permute(0,3,1,2)`
    # (tiny.py:198) broadcasts a (B,H0,W0) tensor against (B,2,H0,W0), which is only well-formed for B == 1
    # (B == 2 silently mixes batch elements with coordinate channels, other B raise), so B == 1 is the
    # behaviour that defines the function.
    pe = torch.cat([tiny.pos_embed(cv[i:i + 1]) for i in range(cv.shape[0])])
    save("tiny", feat0=fa, feat1=fb, corr_volume=cv, pos_embed=pe)


if __name__ == "__main__":
    main()
