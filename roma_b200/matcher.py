"""RegressionMatcher: the public matching API of the reference (romatch/models/matcher.py:425-766) on top of the
sm_100a kernels.  Same method names, argument meaning, return shapes and mutable attributes
(`w_resized, h_resized, upsample_res, upsample_preds, symmetric, attenuate_cert, sample_mode, sample_thresh`).

Deliberate differences (all documented in DESIGN.md):
  * tensor / batched inputs work (the reference raises on them) — uint8 or float (B,3,H,W);
  * no torch.cuda.empty_cache() between passes (matcher.py:666), no second image decode (:670-672);
  * the epilogue (:656-662, :684-730) is one kernel without the `.any()` host sync of :708;
  * there is no CPU path: `device` must be a CUDA device.
"""
from __future__ import annotations

import math
import os
from typing import Optional, Tuple, Union

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F
from PIL import Image

from . import ops

_MEAN = (0.485, 0.456, 0.406)
_STD = (0.229, 0.224, 0.225)


def _load_rgb(im) -> Image.Image:
    if isinstance(im, (str, os.PathLike)):
        im = Image.open(im)
        if im.mode == "I;16":                       # utils.py:656-658
            raise NotImplementedError("Can't handle 16 bit images")
        return im.convert("RGB")
    if isinstance(im, Image.Image):
        if im.mode != "RGB":                        # utils.py:660-662
            raise NotImplementedError("Can't handle non-RGB images")
        return im
    raise TypeError(f"unsupported image input {type(im)}")


def _pil_to_tensor(im: Image.Image, size_hw: Tuple[int, int]) -> torch.Tensor:
    """PIL bicubic resize + /255 (utils.py:165-174, 234-239); normalisation happens on the device."""
    im = im.resize((size_hw[1], size_hw[0]), resample=Image.BICUBIC)
    return torch.from_numpy(np.asarray(im, dtype=np.float32).transpose(2, 0, 1) / 255.0)


class RegressionMatcher(nn.Module):
    def __init__(self, encoder: nn.Module, decoder: nn.Module, h: int = 448, w: int = 448,
                 sample_mode: str = "threshold_balanced", upsample_preds: bool = False, symmetric: bool = False,
                 name: Optional[str] = None, attenuate_cert: Optional[bool] = None):
        super().__init__()
        self.attenuate_cert = attenuate_cert
        self.encoder = encoder
        self.decoder = decoder
        self.name = name
        self.w_resized = w
        self.h_resized = h
        self.sample_mode = sample_mode
        self.upsample_preds = upsample_preds
        self.upsample_res = (14 * 16 * 6, 14 * 16 * 6)
        self.symmetric = symmetric
        self.sample_thresh = 0.05
        self.overlap_encoder = True      # upsample-pass encoder concurrent with the coarse decoder (second stream)
        self.decoder_priority = True     # ... with the coarse decoder on a high-priority stream
        self.native_image_prep = True    # uint8 tensors: normalise + antialiased bicubic resize in one kernel
        self.sample_lanes = 2            # streams sample_batch spreads the pairs over
        self._streams = {}

    # ------------------------------------------------------------------------------------------ helpers
    def _side_stream(self, device) -> "torch.cuda.Stream":
        key = torch.device(device).index if torch.device(device).index is not None else torch.cuda.current_device()
        if key not in self._streams:
            self._streams[key] = torch.cuda.Stream(device=device)
        return self._streams[key]

    def _priority_stream(self, device) -> "torch.cuda.Stream":
        key = ("hi", torch.device(device).index if torch.device(device).index is not None else torch.cuda.current_device())
        if key not in self._streams:
            self._streams[key] = torch.cuda.Stream(device=device, priority=-1)
        return self._streams[key]

    def _second_stream(self, device) -> "torch.cuda.Stream":
        key = ("b", torch.device(device).index if torch.device(device).index is not None else torch.cuda.current_device())
        if key not in self._streams:
            self._streams[key] = torch.cuda.Stream(device=device)
        return self._streams[key]

    def get_output_resolution(self):
        return (self.h_resized, self.w_resized) if not self.upsample_preds else self.upsample_res

    def _normalise(self, im: torch.Tensor) -> torch.Tensor:
        if im.dtype == torch.uint8:
            im = im.float() / 255.0
        mean = torch.tensor(_MEAN, device=im.device, dtype=im.dtype).view(1, 3, 1, 1)
        std = torch.tensor(_STD, device=im.device, dtype=im.dtype).view(1, 3, 1, 1)
        return (im[:, :3] - mean) / std

    def extract_backbone_features(self, batch, batched: bool = True, upsample: bool = False):
        x = torch.cat((batch["im_A"], batch["im_B"]), dim=0)
        return self.encoder(x, upsample=upsample)

    def forward(self, batch, batched: bool = True, upsample: bool = False, scale_factor: float = 1):
        pyr = batch.get("pyramid") or self.extract_backbone_features(batch, upsample=upsample)
        f_q = {s: f.chunk(2)[0] for s, f in pyr.items()}
        f_s = {s: f.chunk(2)[1] for s, f in pyr.items()}
        return self.decoder(f_q, f_s, upsample=upsample, **(batch.get("corresps") or {}), scale_factor=scale_factor)

    def forward_symmetric(self, batch, batched: bool = True, upsample: bool = False, scale_factor: float = 1):
        pyr = batch.get("pyramid") or self.extract_backbone_features(batch, upsample=upsample)
        # support pyramid = query pyramid with the batch halves exchanged (matcher.py:459-463): the decoder swaps
        # after its projections instead of being handed a second, copied pyramid
        return self.decoder(pyr, None, upsample=upsample, **(batch.get("corresps") or {}), scale_factor=scale_factor,
                            swapped=True)

    # ------------------------------------------------------------------------------------------ match
    @torch.inference_mode()
    def match(self, im_A_input, im_B_input, *args, batched: bool = False, device=None, return_logits: bool = False):
        """Dense warp and certainty for an image pair or a batch of pairs (matcher.py:593-730).

        Inputs: two paths, two PIL images (one pair), or two tensors (B,3,H,W) / (3,H,W).  Tensors are either
        uint8 RGB, or float already ImageNet-normalised (what the reference's batched mode expects,
        matcher.py:639-645); any resolution — they are resized on the device (bicubic, antialiased) to the coarse
        and upsample resolutions where they differ.

        Returns (warp, certainty): (H, 2W, 4) and (H, 2W) for one pair, with a leading batch dimension when
        `batched=True`; W doubles only for symmetric models.  `return_logits=True` appends the certainty logits."""
        if device is None:
            device = torch.device("cuda")
        device = torch.device(device)
        if device.type != "cuda":
            raise RuntimeError("roma_b200 runs on sm_100a GPUs only; there is no CPU path (use the reference for CPU)")
        self.train(False)
        hs, ws = self.h_resized, self.w_resized
        tensor_in = torch.is_tensor(im_A_input)
        if tensor_in:
            imA, imB = im_A_input.to(device, non_blocking=True), im_B_input.to(device, non_blocking=True)
            if imA.dim() == 3:
                imA, imB = imA[None], imB[None]
            assert imA.shape == imB.shape, "For batched images we assume same size"
            full_A, full_B = imA, imB
            if imA.dtype == torch.uint8 and self.native_image_prep and imA.shape[1] == 3:
                # (f-3): normalisation + antialiased bicubic resize in one kernel per pass resolution, from the uint8 pixels
                def at(res, t):
                    if max(t.shape[-2] / res[0], t.shape[-1] / res[1]) > 4:          # beyond the kernel's tap budget
                        return F.interpolate(self._normalise(t), size=tuple(res), mode="bicubic", antialias=True, align_corners=False)
                    return ops.prepare_images(t, res)
            else:
                if imA.dtype == torch.uint8:
                    full_A, full_B = self._normalise(imA), self._normalise(imB)

                def at(res, t):
                    return t if tuple(t.shape[-2:]) == tuple(res) else F.interpolate(t, size=tuple(res), mode="bicubic",
                                                                                     antialias=True, align_corners=False)
            coarse = {"im_A": at((hs, ws), full_A), "im_B": at((hs, ws), full_B)}
        else:
            pilA, pilB = _load_rgb(im_A_input), _load_rgb(im_B_input)
            # (f-3): the decoded 8-bit pixels go to the device once; PIL's bicubic resize (bit-exact), / 255 and the
            # normalisation run there in one kernel per pass resolution.  Beyond the kernel's tap budget: PIL on the host.
            devA = torch.from_numpy(np.array(pilA)).to(device, non_blocking=True) if self.native_image_prep else None
            devB = torch.from_numpy(np.array(pilB)).to(device, non_blocking=True) if self.native_image_prep else None

            def at_pil(res, pil, dev_u8):
                if dev_u8 is None or max(pil.size[1] / res[0], pil.size[0] / res[1]) > 4:
                    return self._normalise(_pil_to_tensor(pil, res)[None].to(device))
                return ops.prepare_images_pil(dev_u8, res, channels_last=True)
            coarse = {"im_A": at_pil((hs, ws), pilA, devA), "im_B": at_pil((hs, ws), pilB, devB)}
        symmetric = self.symmetric
        fine, side, corresps = None, None, None
        if self.upsample_preds:
            hs, ws = self.upsample_res
            if tensor_in:
                fine = {"im_A": at((hs, ws), full_A), "im_B": at((hs, ws), full_B)}
            else:
                fine = {"im_A": at_pil((hs, ws), pilA, devA), "im_B": at_pil((hs, ws), pilB, devB)}
            if self.overlap_encoder:
                # The VGG pyramid of the upsample pass does not depend on the coarse pass: run the two concurrently,
                # started once the coarse ENCODER is enqueued, so its large convolutions fill the SMs the
                # latency-bound coarse DECODER (GP solve chain, 40x40 / 70x70 refiners, 1600-token transformer)
                # leaves idle.  Joined before the upsample decoder.
                coarse["pyramid"] = self.extract_backbone_features(coarse, upsample=False)
                cur = torch.cuda.current_stream(device)
                if self.decoder_priority:
                    # the pyramid stays on the caller's stream (enqueued first: few, long kernels) and the coarse
                    # decoder goes to a HIGH-PRIORITY stream: whenever thread blocks of a convolution retire, the
                    # decoder's short dependent kernels are scheduled ahead of the convolution's next blocks
                    hi = self._priority_stream(device)
                    hi.wait_stream(cur)
                    fine["pyramid"] = self.extract_backbone_features(fine, upsample=True)
                    with torch.cuda.stream(hi):
                        corresps = self.forward_symmetric(coarse) if symmetric else self.forward(coarse)
                    cur.wait_stream(hi)
                    for sc in corresps.values():
                        for t in sc.values():
                            if torch.is_tensor(t):
                                t.record_stream(cur)   # allocated on the priority stream, consumed (and freed) on this one
                else:
                    side = self._side_stream(device)
                    side.wait_stream(cur)
                    with torch.cuda.stream(side):
                        fine["pyramid"] = self.extract_backbone_features(fine, upsample=True)
        if corresps is None:
            corresps = self.forward_symmetric(coarse) if symmetric else self.forward(coarse)
        cert16 = corresps[16]["certainty"] if self.attenuate_cert else None
        if self.upsample_preds:
            if side is not None:
                cur.wait_stream(side)
                for t in fine["pyramid"].values():
                    t.record_stream(cur)               # allocated on the side stream, consumed (and freed) on this one
            scale_factor = math.sqrt(self.upsample_res[0] * self.upsample_res[1] / (self.w_resized * self.h_resized))
            fine["corresps"] = corresps[1]
            corresps = (self.forward_symmetric(fine, upsample=True, scale_factor=scale_factor) if symmetric
                        else self.forward(fine, upsample=True, scale_factor=scale_factor))
        out = ops.match_epilogue(corresps[1]["flow"], corresps[1]["certainty"], cert16, symmetric,
                                 return_logits=return_logits)                                          # K3a
        if batched or (tensor_in and im_A_input.dim() == 4):
            return out
        return tuple(o[0] for o in out)

    # ------------------------------------------------------------------------------------------ sample
    def sample(self, matches, certainty, num: int = 10000, generator: Optional[torch.Generator] = None):
        """matcher.py:468-495: threshold -> multinomial(4*num) -> kde -> multinomial(num)."""
        return ops.sample(matches, certainty, num=num, sample_thresh=self.sample_thresh,
                          sample_mode=self.sample_mode, generator=generator)

    def sample_batch(self, matches, certainty, num: int = 10000, generators=None):
        """`sample` for every pair of a batched `match` result: matches (B,H,W,4), certainty (B,H,W) ->
        (list of (num,4), list of (num,)).  Same arithmetic and RNG consumption per pair as B calls of `sample`
        (Philox offsets are assigned at enqueue time, in pair order); consecutive pairs alternate between two
        streams so that one pair's latency-bound multinomial / top-k kernels run under another pair's kde."""
        B = matches.shape[0]
        dev = matches.device
        cur = torch.cuda.current_stream(dev)
        lanes = [self._side_stream(dev), self._second_stream(dev)]
        for k in range(2, max(2, int(self.sample_lanes))):
            key = ("lane", k, dev.index if dev.index is not None else torch.cuda.current_device())
            if key not in self._streams:
                self._streams[key] = torch.cuda.Stream(device=dev)
            lanes.append(self._streams[key])
        for st in lanes:
            st.wait_stream(cur)
        ms, cs = [], []
        for i in range(B):
            with torch.cuda.stream(lanes[i % len(lanes)]):
                m, c = ops.sample(matches[i], certainty[i], num=num, sample_thresh=self.sample_thresh,
                                  sample_mode=self.sample_mode, generator=None if generators is None else generators[i])
            ms.append(m)
            cs.append(c)
        for st in lanes:
            cur.wait_stream(st)
        for t in ms + cs:
            t.record_stream(cur)
        return ms, cs

    # ------------------------------------------------------------------------------------------ keypoints
    def match_keypoints(self, x_A, x_B, warp, certainty, return_tuple: bool = True, return_inds: bool = False):
        """matcher.py:576-591."""
        inds_A, inds_B = ops.match_keypoints(x_A, x_B, warp, certainty, self.sample_thresh)
        if return_tuple:
            return (inds_A, inds_B) if return_inds else (x_A[inds_A], x_B[inds_B])
        return torch.cat((inds_A, inds_B), dim=-1) if return_inds else torch.cat((x_A[inds_A], x_B[inds_B]), dim=-1)

    def conf_from_fb_consistency(self, flow_forward, flow_backward, th: float = 2):
        """matcher.py:530-551: forward-backward consistency mask; flows are (..., H, W, 2)."""
        has_batch = flow_forward.dim() == 4
        if not has_batch:
            flow_forward, flow_backward = flow_forward[None], flow_backward[None]
        H, W = flow_forward.shape[-3:-1]
        th_n = 2 * th / max(H, W)
        dev = flow_forward.device
        xs = torch.linspace(-1 + 1 / W, 1 - 1 / W, W, device=dev)
        ys = torch.linspace(-1 + 1 / H, 1 - 1 / H, H, device=dev)
        coords = torch.stack((xs[None, :].expand(H, W), ys[:, None].expand(H, W)), dim=-1)
        fb = ops.grid_sample(flow_backward.permute(0, 3, 1, 2).float(), flow_forward.float()).permute(0, 2, 3, 1)
        in_th = ((coords - fb).norm(dim=-1) < th_n).float()
        return in_th if has_batch else in_th[0]

    def visualize_warp(self, warp, certainty, im_A=None, im_B=None, im_A_path=None, im_B_path=None, device="cuda",
                       symmetric: bool = True, save_path=None, unnormalize: bool = False):
        """matcher.py:732-766, headless: each image is transferred through the other direction's warp (the K2 gather
        kernel in gather-only mode), blended with white by the certainty; returns the (3, H, W') tensor and, when
        `save_path` is given, writes it as an image file (no window / display involved)."""
        H, W2, _ = warp.shape
        W = W2 // 2 if symmetric else W2
        dev = warp.device
        if im_A is None:
            im_A, im_B = _load_rgb(im_A_path), _load_rgb(im_B_path)
        if not torch.is_tensor(im_A):
            to_t = lambda im: torch.from_numpy(np.asarray(im.resize((W, H)), dtype=np.float32) / 255.0).to(dev).permute(2, 0, 1)
            x_A, x_B = to_t(im_A), to_t(im_B)
        else:
            x_A, x_B = im_A.to(dev).float(), im_B.to(dev).float()
        a_from_b = ops.grid_sample(x_B[None], warp[None, :, :W, 2:].float())[0]
        if symmetric:
            b_from_a = ops.grid_sample(x_A[None], warp[None, :, W:, :2].float())[0]
            warp_im = torch.cat((a_from_b, b_from_a), dim=2)
        else:
            warp_im = a_from_b
        vis = certainty * warp_im + (1 - certainty) * torch.ones_like(certainty)
        if save_path is not None:
            v = vis
            if unnormalize:
                v = v * torch.tensor(_STD, device=dev).view(3, 1, 1) + torch.tensor(_MEAN, device=dev).view(3, 1, 1)
            Image.fromarray((v.clamp(0, 1).permute(1, 2, 0) * 255).round().to(torch.uint8).cpu().numpy()).save(save_path)
        return vis

    # ------------------------------------------------------------------------------------------ coordinates
    def to_pixel_coordinates(self, coords, H_A, W_A, H_B=None, W_B=None):
        if coords.shape[-1] == 2:
            return self._to_pixel_coordinates(coords, H_A, W_A)
        if isinstance(coords, (list, tuple)):
            kA, kB = coords[0], coords[1]
        else:
            kA, kB = coords[..., :2], coords[..., 2:]
        return self._to_pixel_coordinates(kA, H_A, W_A), self._to_pixel_coordinates(kB, H_B, W_B)

    @staticmethod
    def _to_pixel_coordinates(coords, H, W):
        return torch.stack((W / 2 * (coords[..., 0] + 1), H / 2 * (coords[..., 1] + 1)), dim=-1)

    @staticmethod
    def to_normalized_coordinates(coords, H_A, W_A, H_B, W_B):
        if isinstance(coords, (list, tuple)):
            kA, kB = coords[0], coords[1]
        else:
            kA, kB = coords[..., :2], coords[..., 2:]
        kA = torch.stack((2 / W_A * kA[..., 0] - 1, 2 / H_A * kA[..., 1] - 1), dim=-1)
        kB = torch.stack((2 / W_B * kB[..., 0] - 1, 2 / H_B * kB[..., 1] - 1), dim=-1)
        return kA, kB
