"""ctypes wrapper of the plain-C oracle (``oracle/cf_oracle.c``) -- test infrastructure only."""
import ctypes
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "_build", "libcf_oracle.so")
_lib = None


def build(force=False):
    if force or not os.path.exists(LIB_PATH) or os.path.getmtime(LIB_PATH) < os.path.getmtime(os.path.join(HERE, "cf_oracle.c")):
        subprocess.check_call(["make", "-s", "-C", HERE, "-B"])
    return LIB_PATH


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = ctypes.CDLL(LIB_PATH)
        _lib.orc_version.restype = ctypes.c_int
    return _lib


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def pyramid_sizes(h, w, num_levels):
    out = [(h, w)]
    for _ in range(num_levels - 1):
        h, w = h // 2, w // 2
        out.append((h, w))
    return out


def corr_volume(f1, f2):
    f1, f2 = _f32(f1), _f32(f2)
    b, c, h, w = f1.shape
    out = np.empty((b * h * w, 1, h, w), np.float32)
    lib().orc_corr_volume(_p(f1), _p(f2), b, c, h * w, _p(out))
    return out


def avg_pool2x2(x):
    x = _f32(x)
    n, h, w = int(np.prod(x.shape[:-2])), x.shape[-2], x.shape[-1]
    out = np.empty(x.shape[:-2] + (h // 2, w // 2), np.float32)
    lib().orc_avg_pool2x2(_p(x), ctypes.c_long(n), h, w, _p(out))
    return out


def build_pyramid(f1, f2, num_levels=4):
    pyr = [corr_volume(f1, f2)]
    for _ in range(num_levels - 1):
        pyr.append(avg_pool2x2(pyr[-1]))
    return pyr


def lookup(pyramid, coords, radius=4, want_taps=False):
    coords = _f32(coords)
    b, _, h, w = coords.shape
    levels = [_f32(p) for p in pyramid]
    nl, side = len(levels), 2 * radius + 1
    ptrs = (ctypes.c_void_p * nl)(*[p.ctypes.data for p in levels])
    out = np.empty((b, nl * side * side, h, w), np.float32)
    taps = idx = None
    if want_taps:
        taps = np.empty((nl, b * h * w, 2, side), np.float32)
        idx = np.empty((nl, b * h * w, 2, side), np.int32)
    lib().orc_lookup(ptrs, _p(coords), b, h, w, nl, radius, _p(out),
                     _p(taps) if want_taps else None, _p(idx) if want_taps else None)
    return (out, taps, idx) if want_taps else out


def warp2d(flow, src, mode="bilinear", want_taps=False):
    flow, src = _f32(flow), _f32(src)
    b, c, h, w = src.shape
    out = np.empty_like(src)
    taps = np.empty((b, 2, h, w), np.float32) if want_taps else None
    lib().orc_warp2d(_p(flow), _p(src), b, c, h, w, 0 if mode == "bilinear" else 1, _p(out),
                     _p(taps) if want_taps else None)
    return (out, taps) if want_taps else out


def warp_onehot_argmax(flow, label, num_classes=4):
    flow = _f32(flow)
    label = np.ascontiguousarray(label, dtype=np.int64)
    b, _, h, w = label.shape
    out = np.empty((b, 1, h, w), np.int64)
    lib().orc_warp_onehot_argmax(_p(flow), _p(label), num_classes, b, h, w, _p(out))
    return out
