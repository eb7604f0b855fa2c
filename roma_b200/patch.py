"""patch(model): rebind the hot ops of an ALREADY-BUILT reference `romatch` model to the sm_100a kernels.

The seams are module-level names the reference imports once and bound `forward`s:
  romatch.models.matcher.local_correlation   (matcher.py:13, called :124)
  romatch.models.matcher.kde / romatch.models.tiny.kde   (matcher.py:15 / tiny.py:13, called :489 / :321)
  romatch.models.matcher.cls_to_flow_refine  (matcher.py:14, called :383)
  GP.forward of model.decoder.gps[*]         (matcher.py:249, called :377)
  TinyRoMa.corr_volume + pos_embed           (tiny.py:241, :178) -> fused kernel
The reference keeps its own conv bodies, encoders, glue and `match()`; only these calls change.  For the full
fused path (single-launch refiner input assembly, fused update/resample, epilogue kernel, prepared refiner
body) build the model with `roma_b200.roma_outdoor` instead — same state_dict.
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
    if hasattr(model, "corr_volume") and hasattr(model, "pos_embed"):
        model.corr_volume = types.MethodType(_tiny_corr_volume, model)
        model.pos_embed = types.MethodType(_tiny_pos_embed, model)
    return model
