"""ctypes binding of libcardioflow.so (include/cardioflow.h).  No fallback: a missing library is an error."""
import ctypes
import os
import re

import torch

PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(PKG, "libcardioflow.so")
HEADER = os.path.join(os.path.dirname(PKG), "include", "cardioflow.h")

CF_OK, CF_ERR_BAD_ARG, CF_ERR_UNSUPPORTED, CF_ERR_CUDA = 0, -1, -2, -3
PREC_BF16, PREC_FP32, PREC_AUTO, PREC_BF16X3 = 0, 1, 2, 3
DTYPE_F32, DTYPE_BF16 = 0, 1
LAYOUT_NCHW, LAYOUT_NHWC = 0, 1
MODE_BILINEAR, MODE_NEAREST = 0, 1
BUILD_ALLOW_TILED = 1
PRECISIONS = {"bf16": PREC_BF16, "fp32": PREC_FP32, "auto": PREC_AUTO, "bf16x3": PREC_BF16X3}
PRECISION_NAMES = {PREC_BF16: "bf16", PREC_FP32: "fp32", PREC_BF16X3: "bf16x3"}

_c = ctypes
_P, _I, _Z = _c.c_void_p, _c.c_int, _c.c_size_t
_SIGNATURES = {
    "cf_version": (_I, []),
    "cf_last_error": (_c.c_char_p, []),
    "cf_pyramid_level_dims": (_I, [_I, _I, _I, _c.POINTER(_I), _c.POINTER(_I)]),
    "cf_corr_pyramid_level_offset": (_Z, [_I, _I, _I, _I]),
    "cf_corr_pyramid_bytes": (_Z, [_I, _I, _I, _I]),
    "cf_coords_grid": (_I, [_P, _I, _I, _I, _P]),
    "cf_corr_build_workspace_bytes": (_Z, [_I, _I, _I, _I, _I]),
    "cf_corr_build": (_I, [_P, _P, _I, _I, _I, _I, _I, _I, _P, _P, _Z, _c.POINTER(_I), _P]),
    "cf_corr_build_layout": (_I, [_P, _P, _I, _I, _I, _I, _I, _I, _I, _P, _P, _Z, _c.POINTER(_I), _c.POINTER(_c.c_uint), _P]),
    "cf_corr_build_ex": (_I, [_P, _P, _I, _I, _I, _I, _I, _I, _I, _I, _I, _P, _P, _Z, _c.POINTER(_I), _c.POINTER(_c.c_uint), _P]),
    "cf_corr_lookup_layout": (_I, [_P, _P, _I, _I, _I, _I, _I, _c.c_uint, _P, _P]),
    "cf_corr_pool": (_I, [_P, _I, _I, _I, _I, _I, _P]),
    "cf_corr_lookup": (_I, [_P, _P, _I, _I, _I, _I, _I, _P, _P]),
    "cf_corr_lookup_taps": (_I, [_P, _I, _I, _I, _I, _I, _P, _P, _P]),
    "cf_fmap_pyramid_level_offset": (_Z, [_I, _I, _I, _I, _I]),
    "cf_fmap_pyramid_bytes": (_Z, [_I, _I, _I, _I, _I]),
    "cf_fmap_pyramid_build": (_I, [_P, _I, _I, _I, _I, _I, _P, _P]),
    "cf_corr_lookup_fused": (_I, [_P, _P, _P, _I, _I, _I, _I, _I, _I, _P, _P]),
    "cf_corr_otf_supported": (_I, [_I, _I, _I, _I, _I]),
    "cf_corr_otf_operand_bytes": (_Z, [_I, _I, _I, _I, _I]),
    "cf_corr_otf_prepare": (_I, [_P, _P, _I, _I, _I, _I, _I, _P, _P, _Z, _P]),
    "cf_corr_lookup_otf": (_I, [_P, _P, _I, _I, _I, _I, _I, _I, _P, _P]),
    "cf_warp2d": (_I, [_P, _P, _I, _I, _I, _I, _I, _P, _P]),
    "cf_warp2d_taps": (_I, [_P, _I, _I, _I, _P, _P]),
    "cf_warp2d_onehot_argmax": (_I, [_P, _P, _I, _I, _I, _I, _P, _P]),
    "cf_warp2d_onehot_argmax_counts": (_I, [_P, _P, _I, _I, _I, _I, _P, _P, _P]),
    "cf_warp2d_onehot_argmax_u8": (_I, [_P, _P, _I, _I, _I, _I, _P, _P, _I, _P]),
    "cf_warp2d_bwd": (_I, [_P, _P, _P, _I, _I, _I, _I, _P, _P, _P]),
    "cf_warp2d_nearest_bwd": (_I, [_P, _P, _I, _I, _I, _I, _P, _P]),
    "cf_corr_lookup_bwd": (_I, [_P, _P, _I, _I, _I, _I, _I, _P, _P]),
    "cf_corr_pool_bwd": (_I, [_P, _I, _I, _I, _I, _P]),
    "cf_corr_volume_bwd": (_I, [_P, _P, _P, _I, _I, _I, _I, _P, _P, _P]),
    "cf_corr_volume_bwd_workspace_bytes": (_Z, [_I, _I, _I, _I]),
    "cf_corr_volume_bwd_noalloc": (_I, [_P, _P, _P, _I, _I, _I, _I, _P, _P, _P]),
    "cf_corr_volume_bwd_ex": (_I, [_P, _P, _P, _I, _I, _I, _I, _I, _P, _P, _P, _Z, _c.POINTER(_I), _P]),
    "cf_vecint": (_I, [_P, _I, _I, _I, _I, _P, _P, _P]),
    "cf_sample_points": (_I, [_P, _P, _I, _I, _I, _I, _I, _I, _I, _I, _P, _P]),
    "cf_sample_points_bwd": (_I, [_P, _P, _P, _I, _I, _I, _I, _I, _I, _I, _I, _P, _P, _P]),
    "cf_local_corr": (_I, [_P, _P, _I, _I, _I, _I, _I, _I, _c.c_float, _I, _P, _P]),
    "cf_local_corr_bwd": (_I, [_P, _P, _P, _I, _I, _I, _I, _I, _I, _c.c_float, _I, _P, _P, _P]),
    "cf_warp2d_xy": (_I, [_P, _P, _I, _I, _I, _I, _P, _P]),
    "cf_warp2d_xy_bwd": (_I, [_P, _P, _P, _I, _I, _I, _I, _P, _P, _P]),
    "cf_upsample_convex": (_I, [_P, _P, _I, _I, _I, _I, _I, _c.c_float, _P, _P]),
    "cf_upsample_convex_bwd": (_I, [_P, _P, _P, _I, _I, _I, _I, _I, _c.c_float, _P, _P, _P]),
    "cf_sum_f64_workspace_bytes": (_Z, []),
    "cf_sum_f64": (_I, [_P, _Z, _P, _P, _Z, _P]),
    "cf_ncc_workspace_bytes": (_Z, [_I, _I, _I, _I]),
    "cf_ncc": (_I, [_P, _P, _I, _I, _I, _I, _c.c_float, _P, _P, _P, _Z, _P]),
    "cf_ncc_bwd": (_I, [_P, _P, _I, _I, _I, _I, _c.c_float, _P, _I, _c.c_float, _P, _P, _P, _Z, _P]),
}

_lib = None


def declared_symbols():
    """Every entry point include/cardioflow.h declares (used by the symbol-export test)."""
    with open(HEADER) as f:
        return sorted(set(re.findall(r"^CF_API[^\n(]*?\b(cf_\w+)\(", f.read(), flags=re.M)))


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'`. "
                "cardioflow_b200 has no CPU / eager fallback.")
        handle = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(handle, name)
            fn.restype, fn.argtypes = res, args
        _lib = handle
    return _lib


def check(rc, what=""):
    if rc == CF_OK:
        return
    msg = lib().cf_last_error().decode() or f"{what} failed with code {rc}"
    if rc in (CF_ERR_BAD_ARG, CF_ERR_UNSUPPORTED):
        raise ValueError(msg)
    raise RuntimeError(msg)


def stream():
    return torch.cuda.current_stream().cuda_stream


def require_cuda(*tensors):
    for t in tensors:
        if not t.is_cuda:
            raise RuntimeError("cardioflow_b200 runs on CUDA tensors only (there is no CPU fallback; "
                               "the CPU restatement lives in oracle/ and is test infrastructure)")
    dev = tensors[0].device
    for t in tensors[1:]:
        if t.device != dev:
            raise ValueError(f"tensors on different devices: {dev} vs {t.device}")
    return dev


def f32c(t):
    """fp32 contiguous view/copy (RAFT's CorrBlock does .float() on its inputs as well)."""
    return t.detach().to(torch.float32).contiguous() if (t.dtype != torch.float32 or not t.is_contiguous()) else t.detach()
