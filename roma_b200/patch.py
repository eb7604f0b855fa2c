"""patch(model): rebind the hot ops of an ALREADY-BUILT reference `romatch` model to the sm_100a kernels.

The seams are module-level names the reference imports once and bound `forward`s:
  romatch.models.matcher.local_correlation   (matcher.py:13, called :124)
  romatch.models.matcher.kde / romatch.models.tiny.kde   (matcher.py:15 / tiny.py:13, called :489 / :321)
  romatch.models.matcher.cls_to_flow_refine  (matcher.py:14, called :383)
  GP.forward of model.decoder.gps[*]         (matcher.py:249, called :377)
  TinyRoMa.corr_volume + pos_embed           (tiny.py:241, :178) -> fused kernel
  ConvRefiner.forward of model.decoder.conv_refiner[*]   (matcher.py:105-143, called :393-395): grid_sample + disp_emb +
      local_correlation + cat in ONE launch (roma_refiner_gather); the reference's own block1 / hidden_blocks / out_conv follow
  model.sample / model.match_keypoints       (matcher.py:468-495 / :576-591, tiny.py:297-327): same rows / indices
The reference keeps its own conv bodies, encoders, `Decoder.forward` glue (update / resample, :397-417) and `match()`
with its epilogue (:656-730); only the calls above change.  Those two — and the prepared refiner body, the fused
update/resample and epilogue kernels, tensor / batched inputs — need the factory path: build the model with
`roma_b200.roma_outdoor` instead — same state_dict.
"""
from __future__ import annotations

import types

import torch

from . import ops


def _gp_forward(self, x, y, **kwargs):
    return ops.gp_posterior(x, y, self.pos_conv.weight, self.pos_conv.bias, T=float(self.K.T),
                            sigma_noise=float(self.sigma_noise))


def _cls_to_flow_refine(cls):
    return ops.cls_to_flow_refine(cls)


class _FusedVolume:
    """Stand-in for the (B,H1,W1,H0,W0) volume: carries the features so pos_embed can run the fused kernel."""

    def __init__(self, feat0, feat1):
        self.feat0, self.feat1 = feat0, feat1


def _tiny_corr_volume(self, feat0, feat1):
    return _FusedVolume(feat0, feat1)


def _tiny_pos_embed(self, corr_volume):
    if isinstance(corr_volume, _FusedVolume):
        if self.training or getattr(self, "exact_softmax", False):
            raise NotImplementedError("the fused Tiny kernel implements the inference branch of pos_embed (tiny.py:187-199)")
        return ops.tiny_corr_posembed(corr_volume.feat0, corr_volume.feat1)
    return type(self).pos_embed(self, corr_volume)


def _refiner_forward(self, x, y, flow, scale_factor=1, logits=None):
    """ConvRefiner.forward (matcher.py:105-143) with lines 109-132 as one kernel launch; the conv stack stays the reference's."""
    from torch.nn import functional as F  # noqa: F401
    if self.concat_logits or getattr(self, "no_im_B_fm", False) or self.sample_mode != "bilinear":
        return type(self).forward(self, x, y, flow, scale_factor=scale_factor, logits=logits)
    dt = self.amp_dtype if (self.amp and x.is_cuda) else torch.float32
    x, y = x.to(dt), y.to(dt)                                   # what autocast hands the reference's cat (16-bit proj outputs)
    dw = self.disp_emb.weight if self.has_displacement_emb else None
    db = self.disp_emb.bias if self.has_displacement_emb else None
    r = self.local_corr_radius if self.has_displacement_emb else None
    d = ops.refiner_inputs(x, y, flow, dw, db, 40 / 32 * scale_factor, r)
    with torch.autocast("cuda", enabled=bool(self.amp), dtype=self.amp_dtype):
        d = self.block1(d)
        d = self.hidden_blocks(d)
    d = self.out_conv(d.float())
    return d[:, :-1], d[:, -1:]


def _sample(self, matches, certainty, num=10000):
    from .matcher import RegressionMatcher
    return RegressionMatcher.sample(self, matches, certainty, num=num)


def _tiny_sample(self, matches, certainty, num=5000):
    from .host.tiny import TinyRoMa
    return TinyRoMa.sample(self, matches, certainty, num=num)


def _match_keypoints(self, x_A, x_B, warp, certainty, return_tuple=True, return_inds=False):
    from .matcher import RegressionMatcher
    return RegressionMatcher.match_keypoints(self, x_A, x_B, warp, certainty, return_tuple=return_tuple, return_inds=return_inds)


def patch(model, matcher_module=None, tiny_module=None):
    """Rebind the seams for `model` (a romatch RegressionMatcher or TinyRoMa).  Returns the model.

    `matcher_module` / `tiny_module` default to the imported `romatch.models.matcher` / `romatch.models.tiny`
    (passed explicitly by tests that use stand-ins)."""
    if matcher_module is None or tiny_module is None:
        import importlib
        matcher_module = matcher_module or importlib.import_module("romatch.models.matcher")
        tiny_module = tiny_module or importlib.import_module("romatch.models.tiny")
    matcher_module.local_correlation = ops.local_correlation
    matcher_module.kde = ops.kde
    matcher_module.cls_to_flow_refine = _cls_to_flow_refine
    tiny_module.kde = ops.kde
    decoder = getattr(model, "decoder", None)
    if decoder is not None and hasattr(decoder, "gps"):
        for gp in decoder.gps.values():
            gp.forward = types.MethodType(_gp_forward, gp)
        for ref in getattr(decoder, "conv_refiner", {}).values():
            if hasattr(ref, "block1") and hasattr(ref, "out_conv"):
                ref.forward = types.MethodType(_refiner_forward, ref)
    if hasattr(model, "corr_volume") and hasattr(model, "pos_embed"):
        model.corr_volume = types.MethodType(_tiny_corr_volume, model)
        model.pos_embed = types.MethodType(_tiny_pos_embed, model)
        if hasattr(model, "sample_thresh"):
            model.sample = types.MethodType(_tiny_sample, model)
    elif hasattr(model, "sample_thresh") and hasattr(model, "sample_mode"):
        model.sample = types.MethodType(_sample, model)
        model.match_keypoints = types.MethodType(_match_keypoints, model)
    return model
