"""ctypes binding of libroma_b200.so — the C ABI declared in include/roma_b200.h.

There is deliberately no fallback: if the shared library is missing or an entry point fails, the
caller gets an exception.  (The reference has no FFI at all — SURVEY.md §8b — so these signatures
are the boundary this build defines; INTEGRATION.md shows the same stub from the reference's side.)
"""
from __future__ import annotations

import ctypes as C
from pathlib import Path

LIB_PATH = Path(__file__).resolve().parent / "libroma_b200.so"

ROMA_F32, ROMA_F16, ROMA_BF16 = 0, 1, 2
ABI_VERSION = 1


class Strides4(C.Structure):
    _fields_ = [("b", C.c_int64), ("c", C.c_int64), ("y", C.c_int64), ("x", C.c_int64)]


class RomaError(RuntimeError):
    pass


_vp, _i, _f, _sz = C.c_void_p, C.c_int, C.c_float, C.c_size_t

# name -> (restype, argtypes); must list every symbol include/roma_b200.h declares
SIGNATURES = {
    "roma_abi_version": (_i, []),
    "roma_last_error_string": (C.c_char_p, []),
    "roma_check_device": (_i, [_i]),
    "roma_kernel_launches": (C.c_ulonglong, []),
    "roma_refiner_gather": (_i, [_vp, _vp, _vp, Strides4, _vp, _vp, _f,
                                 _vp, _i, _i, _i, _i, _i,
                                 _i, _i, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "roma_refiner_gather_workspace_bytes": (_sz, [_i, _i, _i, _i]),
    "roma_refiner_gather_ws": (_i, [_vp, _vp, _vp, Strides4, _vp, _vp, _f,
                                    _vp, _i, _i, _i, _i, _i,
                                    _i, _i, _i, _i, _i, _i, _i, _i, _i, _vp, _sz, _vp]),
    "roma_flow_update_resample": (_i, [_vp, Strides4, _vp, Strides4, _vp, Strides4, _f, _f, _f,
                                       _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _vp]),
    "roma_match_epilogue": (_i, [_vp, Strides4, _vp, Strides4, _vp, Strides4, _i, _i,
                                 _vp, _vp, _vp, _i, _i, _i, _i, _vp]),
    "roma_gp_workspace_bytes": (_sz, [_i, _i, _i, _i, _i]),
    "roma_gp_set_path": (_i, [_i]),
    "roma_gp_set_lanes": (_i, [_i]),
    "roma_gp_set_planes": (_i, [_i]),
    "roma_local_corr_set_path": (_i, [_i]),
    "roma_tc_gemm_nt": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _vp]),
    "roma_gp_posterior": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i,
                               _f, _f, _f, _i, _vp, _sz, _vp]),
    "roma_kde_workspace_bytes": (_sz, [_i]),
    "roma_kde_set_direct": (_i, [_i]),
    "roma_kde": (_i, [_vp, _i, _i, C.c_double, _i, _vp, _vp, _vp, _sz, _vp]),
    "roma_kde_debug_d2": (_i, [_vp, _i, _i, _vp, _vp]),
    "roma_kde_debug_lut": (_i, [_vp, C.c_double, _vp]),
    "roma_match_keypoints_workspace_bytes": (_sz, [_i, _i]),
    "roma_match_keypoints": (_i, [_vp, _i, _vp, _i, _vp, _vp, _i, _i, _f, _vp, _vp, _i, _vp, _vp, _sz, _vp]),
    "roma_match_keypoints_static_max_points": (_i, []),
    "roma_match_keypoints_static": (_i, [_vp, _i, _vp, _i, _vp, _vp, _i, _i, _f, _vp, _vp, _i, _vp, _vp]),
    "roma_dwconv5x5_bias_act": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _vp]),
    "roma_dwconv5x5_pointwise": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _vp]),
    "roma_dwconv_set_path": (_i, [_i]),
    "roma_tc_set_tma": (_i, [_i]),
    "roma_pointwise_narrow_set_simt": (_i, [_i]),
    "roma_layernorm": (_i, [_vp, _vp, _vp, _vp, C.c_longlong, _i, C.c_float, _i, _vp]),
    "roma_layernorm_shift": (_i, [_vp, _vp, _vp, _vp, _vp, C.c_longlong, _i, C.c_float, _i, _vp]),
    "roma_layernorm_cast": (_i, [_vp, _vp, _vp, _vp, C.c_longlong, _i, C.c_float, _i, _vp]),
    "roma_residual_add": (_i, [_vp, _vp, _vp, C.c_longlong, _i, _vp]),
    "roma_add_layernorm": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, C.c_longlong, _i, C.c_float, _i, _vp]),
    "roma_topk_workspace_bytes": (C.c_size_t, [C.c_longlong, C.c_longlong]),
    "roma_topk_desc": (_i, [_vp, _i, C.c_longlong, C.c_longlong, _vp, _vp, _vp, C.c_size_t, _vp]),
    "roma_gelu": (_i, [_vp, _vp, C.c_longlong, _i, _vp]),
    "roma_pointwise_narrow": (_i, [_vp, _vp, _vp, _vp, C.c_longlong, _i, _i, _i, _vp]),
    "roma_maxpool2x2_nhwc": (_i, [_vp, _vp, _i, _i, _i, _i, _i, _vp]),
    "roma_refiner_out_conv": (_i, [_vp, _vp, _vp, _vp, C.c_longlong, _i, _i, _i, _vp]),
    "roma_cls_to_flow_refine": (_i, [_vp, Strides4, _i, _vp, _i, _i, _i, _i, _vp]),
    "roma_prepare_images": (_i, [_vp, _vp, _i, _i, _i, _i, _i, _vp, _vp, _vp]),
    "roma_prepare_images_pil": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _i, C.c_longlong, C.c_longlong, C.c_longlong,
                                     C.c_longlong, _vp, _vp, _vp]),
    "roma_tiny_corr_posembed": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _vp]),
    "roma_tiny_corr_posembed_workspace_bytes": (_sz, [_i, _i, _i, _i, _i]),
    "roma_tiny_corr_posembed_tc": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _vp, _sz, _vp]),
}

_lib = None


def load() -> C.CDLL:
    """Load the library once; raise (never fall back) if it is missing or stale."""
    global _lib
    if _lib is not None:
        return _lib
    if not LIB_PATH.exists():
        raise RomaError(
            f"{LIB_PATH} not found: build it with `python -m roma_b200.build` "
            "(roma_b200 has no CPU or PyTorch fallback for its kernels)")
    lib = C.CDLL(str(LIB_PATH))
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)   # AttributeError if the .so does not export a declared symbol
        fn.restype = res
        fn.argtypes = args
    if lib.roma_abi_version() != ABI_VERSION:
        raise RomaError(f"libroma_b200 ABI {lib.roma_abi_version()} != expected {ABI_VERSION}; rebuild")
    _lib = lib
    return lib


def check(rc: int, what: str) -> None:
    if rc == 0:
        return
    msg = load().roma_last_error_string().decode(errors="replace")
    if rc < 0:
        raise ValueError(f"{what} rejected its arguments (code {rc}): {msg}")
    raise RomaError(f"{what} failed with CUDA error {rc}: {msg}")
