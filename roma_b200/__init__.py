"""roma_b200 — B200-native (sm_100a) implementation of RoMa's dense-matching hot path, behind the romatch API.

    from roma_b200 import roma_outdoor, roma_indoor, tiny_roma_v1_outdoor      # same factories as `romatch`
    model = roma_outdoor(device="cuda", weights=..., dinov2_weights=...)
    warp, certainty = model.match(im_A, im_B)
    matches, cert = model.sample(warp, certainty)

`roma_b200.ops` holds the tensor-level wrappers over the C ABI (include/roma_b200.h); `roma_b200.patch(model)`
rebinds the hot ops of an already-built reference `romatch` model to the kernels.
"""
from .factory import (roma_model, roma_outdoor, roma_indoor, tiny_roma_v1_model, tiny_roma_v1_outdoor,  # noqa: F401
                      random_init_weights, state_digest, weight_urls)
from .matcher import RegressionMatcher  # noqa: F401
from .patch import patch  # noqa: F401

__version__ = "0.1.0"
